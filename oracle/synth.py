"""Re-export of gndnet_b200/synth.py (the synthetic workload generator) for the oracle harness and tests."""
from gndnet_b200.synth import *  # noqa: F401,F403
from gndnet_b200.synth import _CFG, _GOLDEN  # noqa: F401
