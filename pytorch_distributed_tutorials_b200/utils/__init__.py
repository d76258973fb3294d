from .seed import set_random_seeds  # noqa: F401
from .checkpoint import save_checkpoint, load_checkpoint  # noqa: F401
from .timing import StepTimer, ClockSampler  # noqa: F401
