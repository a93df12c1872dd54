"""B200-native association engine behind the BoxMOT tracker API.

Public surface mirrors the reference package root (boxmot/__init__.py:6,25-37):
``create_tracker``, ``get_tracker_config`` and the tracker classes."""
__version__ = "0.1.0"

from boxmot_b200.tracker_zoo import create_tracker, get_tracker_config  # noqa: F401
from boxmot_b200.trackers.boosttrack import BoostTrack  # noqa: F401
from boxmot_b200.trackers.botsort import BotSort  # noqa: F401
from boxmot_b200.trackers.bytetrack import ByteTrack  # noqa: F401
from boxmot_b200.trackers.deepocsort import DeepOcSort  # noqa: F401
from boxmot_b200.trackers.ocsort import OcSort  # noqa: F401

__all__ = ["create_tracker", "get_tracker_config", "OcSort", "ByteTrack", "BotSort", "DeepOcSort", "BoostTrack"]
