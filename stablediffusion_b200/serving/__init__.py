"""Serving-side pieces of the hot path: the output stage of `runpod-worker/handler_logic.py`."""
from .batching import RequestBatcher  # noqa: F401
from .handler_logic import OutputStage, convert_pt_to_numpy  # noqa: F401
