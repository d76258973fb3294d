"""Same launch path as the reference: ``torchrun --nproc_per_node=N resnet/main.py [flags]``."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from pytorch_distributed_tutorials_b200.train import main  # noqa: E402

if __name__ == "__main__":
    main()
