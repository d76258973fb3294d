"""pytorch-distributed-tutorials_b200 — a Blackwell-native (sm_100a) data-parallel training framework with the
capabilities of chkda/pytorch-distributed-tutorials' ``resnet/main.py``.  See DESIGN.md."""
__version__ = "0.1.0"
