"""Seeded (query set, target frame) scenarios for the guided-search tests (generator: synth.synth_guided_case)."""
from orb_slam2_detailed_comments_b200._lib import KP_DTYPE  # noqa: F401
from orb_slam2_detailed_comments_b200.synth import synth_guided_case as make_case  # noqa: F401
