"""B200-native hot path of tim-hua-01/introspective-awareness: concept-vector extraction and
residual-stream injection during batched steered decoding.  See DESIGN.md / INTEGRATION.md."""

from . import grid, model_utils, parallel, steering_utils, sweep, vector_utils  # noqa: F401

__all__ = ["model_utils", "vector_utils", "steering_utils", "grid", "parallel", "sweep"]
