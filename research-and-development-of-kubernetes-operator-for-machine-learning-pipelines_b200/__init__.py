"""B200-native canary metric-validation step for the MLflow->Seldon operator.

Only the hot path named by BASELINE.json's north_star lives here: batched
`get_model_metrics` -> `should_promote_model` -> traffic-weight step loop
(/root/reference/mlflow_operator.py:288-460), as hand-written sm_100a CUDA kernels
behind a C ABI (include/canary_gate.h), with the reference's own Python signatures
on top (operator_api).  There is no CPU fallback: importing the compute modules
without the built CUDA library raises.
"""
from . import params  # noqa: F401  (ctypes mirrors only; safe without the library)

__all__ = ["params"]
