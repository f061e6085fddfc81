"""Re-export of the seeded generators (they live in the product package so bench.py can use them
without touching oracle/)."""
from yolo_series_b200.synth import seeded_state_dict, synth_image, synth_pred  # noqa: F401
