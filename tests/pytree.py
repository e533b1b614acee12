"""moved to pt_b200/pytree.py (bench.py and the host drivers use it); kept as an alias for the tests"""
from pt_b200.pytree import *  # noqa: F401,F403
from pt_b200 import pytree as _m
globals().update({k: v for k, v in vars(_m).items() if not k.startswith("__")})
