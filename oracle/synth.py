"""Seeded synthetic inputs / random checkpoints: re-export of speakerverification_b200.synth (pure data generation,
no arithmetic of the path), so tests and make_golden.py keep importing `oracle.synth`."""
from speakerverification_b200.synth import *  # noqa: F401,F403
