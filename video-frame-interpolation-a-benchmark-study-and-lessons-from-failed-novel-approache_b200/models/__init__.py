"""Mirror of the reference's `models` package for the hot path (base interface, RIFE, adaptive)."""
from .base import BaseModel, InferenceResult, JointModel, ModelInfo, TwoStageModel  # noqa: F401
