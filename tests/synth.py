"""moved to pt_b200/synth.py (bench.py and the host drivers use it); kept as an alias for the tests"""
from pt_b200.synth import *  # noqa: F401,F403
from pt_b200 import synth as _m
globals().update({k: v for k, v in vars(_m).items() if not k.startswith("__")})
