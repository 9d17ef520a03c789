"""Synthetic inputs for tests: re-export of the package's standalone generator (slam-rs_b200/synth.py),
loaded by file path so that importing the oracle never loads liborbx."""
import importlib.util as _u
import os as _os

_p = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))), "slam-rs_b200", "synth.py")
_spec = _u.spec_from_file_location("_orbx_synth", _p)
_m = _u.module_from_spec(_spec)
_spec.loader.exec_module(_m)
frame, sequence, random_descriptors, map_and_frame_descriptors = (
    _m.frame, _m.sequence, _m.random_descriptors, _m.map_and_frame_descriptors)
