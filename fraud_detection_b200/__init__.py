"""fraud_detection_b200 - the batch scoring + SHAP hot path of wtfashwin/fraud-detection,
rebuilt for NVIDIA B200 (sm_100a).

Host side mirrors the reference's duck-typed call surface (``transform``, ``predict``,
``predict_proba``, ``explainer.shap_values``); underneath is ``libfdb200.so`` - a C-ABI
library of hand-written CUDA kernels, bound with ctypes.  Importing this package requires
the built library (``python fraud_detection_b200/build.py``); there is no CPU fallback.

``device`` (torch tensor in / tensor out) and ``distributed`` are imported on demand so the
serving classes do not pull torch into a gunicorn/Celery worker.
"""

from ._lib import FdbError, PinnedBuffer, get_ctx, host_near_gpu, lib  # noqa: F401
from .estimators import B200LogisticRegression, B200StandardScaler, from_joblib, wrap  # noqa: F401
from .explainers import B200KernelExplainer, B200LinearExplainer, KernelShapPlan, KernelShapStream  # noqa: F401

__version__ = "0.1.0"
