"""fod_b200 — B200-native hot path of atonderski/future-object-detection.

Public surface (mirrors the reference): `build_model(args, detr_args)` (runs/_model.py), `SpatioTemporalDETRArgs`
and `SpatioTemporalDETR` (future_od/models/st_detr.py).  Everything computes through libfod_b200.so (C ABI,
include/fod_b200.h); importing the model modules requires torch, running them requires the built library and a B200.
"""
__version__ = "0.2.0"


def __getattr__(name):  # lazy: `import fod_b200` stays cheap and CPU-safe
    if name in ("build_model", "SpatioTemporalDETRArgs", "SpatioTemporalDETR", "HungarianMatcher", "SetCriterion",
                "FuturePredCore", "SeparateEncoder", "JointEncoder", "set_dropout"):
        from . import st_detr
        return getattr(st_detr, name)
    if name in ("ApAccumulator", "aggregate_mean_average_precision"):  # future_od/utils/od_map.py, trainer.py:205-236
        from . import od_map
        return getattr(od_map, name)
    raise AttributeError(name)
