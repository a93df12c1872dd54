"""create_tracker / get_tracker_config with the reference signature
(boxmot/tracker_zoo.py:8-93): YAML ``default`` values become constructor arguments,
unknown tracker names print a message and raise KeyError.

Side-input contract (all six trackers): ReID models and camera-motion estimators are OUTSIDE the
association path.  ``reid_weights`` / ``device`` / ``half`` are accepted and ignored; the trackers that
use appearance (botsort, deepocsort, boosttrack with_reid, strongsort) need the detections'
embeddings as ``update(dets, img, embs)`` -- the reference runs its ReID backend when ``embs`` is
None, these classes raise RuntimeError instead -- and take the 2x3 warp that the reference's
``self.cmc.apply(img, dets)`` would return as the extra keyword ``update(..., warp=H)`` (identity /
no correction when omitted)."""
from pathlib import Path

import yaml

TRACKER_CONFIGS = Path(__file__).resolve().parent / "configs"


def get_tracker_config(tracker_type):
    """Returns the path to the tracker configuration file."""
    return TRACKER_CONFIGS / f"{tracker_type}.yaml"


_MAPPING = {
    "ocsort": ("boxmot_b200.trackers.ocsort", "OcSort"),
    "bytetrack": ("boxmot_b200.trackers.bytetrack", "ByteTrack"),
    "botsort": ("boxmot_b200.trackers.botsort", "BotSort"),
    "deepocsort": ("boxmot_b200.trackers.deepocsort", "DeepOcSort"),
    "boosttrack": ("boxmot_b200.trackers.boosttrack", "BoostTrack"),
    "strongsort": ("boxmot_b200.trackers.strongsort", "StrongSort"),
}
_REID_TRACKERS = ("strongsort", "botsort", "deepocsort", "hybridsort", "boosttrack")


def create_tracker(tracker_type, tracker_config=None, reid_weights=None, device=None, half=None,
                   per_class=None, evolve_param_dict=None):
    if evolve_param_dict is None:
        with open(tracker_config, "r") as f:
            yaml_config = yaml.load(f, Loader=yaml.FullLoader)
            tracker_args = {param: details["default"] for param, details in yaml_config.items()}
    else:
        tracker_args = dict(evolve_param_dict)
    if tracker_type not in _MAPPING:
        print("Error: No such tracker found.")
    module_path, class_name = _MAPPING[tracker_type]
    tracker_class = getattr(__import__(module_path, fromlist=[class_name]), class_name)
    tracker_args["per_class"] = per_class
    if tracker_type in _REID_TRACKERS:
        tracker_args.update({"reid_weights": reid_weights, "device": device, "half": half})
        if tracker_type == "strongsort":
            tracker_args.pop("per_class")
            # M2: strongsort.yaml carries six keys StrongSort.__init__ does not take
            # (the reference raises TypeError here); only the constructor's own arguments are passed
            import inspect
            accepted = inspect.signature(tracker_class.__init__).parameters
            tracker_args = {k: v for k, v in tracker_args.items() if k in accepted}
    return tracker_class(**tracker_args)
