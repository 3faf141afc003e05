"""Minimal stand-ins for the reference's Map / Frame / Point containers (src/v2/map.py, frame.py, point.py), as far as
LocalBA.py touches them: they live next to the synthetic generators (visual_slam_b200/synth.py) because bench.py's
latency block drives the drop-in BundleAdjustment through them too."""
from visual_slam_b200.synth import IclCamera as _Cam  # noqa: F401
from visual_slam_b200.synth import SynthFrame as _Frame  # noqa: F401
from visual_slam_b200.synth import SynthMap as _Map  # noqa: F401
from visual_slam_b200.synth import SynthPoint as _Point  # noqa: F401
from visual_slam_b200.synth import map_from_problem as _map_from_problem  # noqa: F401
