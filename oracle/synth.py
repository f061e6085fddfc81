"""Re-export of the seeded generators (they live in the product package so bench.py can use them
without touching oracle/)."""
from yolo_series_b200.synth import seeded_state_dict, synth_image, synth_pred, synth_pred_wide  # noqa: F401
