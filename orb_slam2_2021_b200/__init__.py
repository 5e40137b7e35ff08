"""orb_slam2_2021_b200 — B200-native ORB front-end (ORBextractor + Frame::ComputeStereoMatches).

The compute lives in liborbx.so (hand-written sm_100a CUDA behind the C-ABI of include/orbx.h); this
package is the thin host-side mirror of the reference interface. Importing the package is cheap and
works without a GPU; any compute call without the built library or without a CUDA device raises.
"""
from ._lib import KP_DTYPE, OrbxError, SO_PATH, build, lib  # noqa: F401
from .extractor import ORBextractor, PinnedArray, host_numa_node, set_tuning, StereoFrameBatch, compute_stereo_matches, compute_stereo_matches_static, compute_stereo_matches_resident, stereo_frame  # noqa: F401
from .frame import FrameView, MapPointView, Matcher, TrackPointView  # noqa: F401
from .bow import ORBVocabulary  # noqa: F401
