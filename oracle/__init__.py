"""CPU oracle for the fraud-detection scoring + SHAP hot path.

TEST INFRASTRUCTURE ONLY.  Nothing in ``fraud_detection_b200`` may import this
package; only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s
``cpu_baseline`` / ``--impl reference`` legs do, and there only as the checker or
the timed CPU baseline - never as the product path.

What it restates (citations relative to the upstream repository; ``SK:`` is the
scikit-learn installed in this image):

* ``scoring``      - StandardScaler.transform / LogisticRegression.decision_function /
                     predict_proba / predict (``api/app.py:194,209,223``;
                     ``SK:preprocessing/_data.py:1092-1135``,
                     ``SK:linear_model/_base.py:366-440``).  PINNED: checked against the
                     live sklearn objects unpickled from the reference's own
                     ``models/*.joblib`` via ``tests/golden/scoring_golden.npz``.
* ``linear_shap``  - shap.LinearExplainer(interventional) as used by
                     ``explain_model.py:24-27,39-40`` and ``api/worker.py:52-53,75`` and the
                     hand-rolled ``coef * x`` attribution of ``xai_tasks.py:106-107``.
* ``kernel_shap``  - shap.KernelExplainer (``shap==0.49.1``, ``requirements.txt:21``).

PARITY UNPINNED for the two shap restatements: the ``shap`` package is not in
this image, not on disk and not installable (no network), and the reference
holds no golden vector, test or fixture for any SHAP value (``tests/test_api.py``
asserts HTTP status codes only).  They follow the published algorithm as
recalled from upstream ``shap/explainers/_linear.py`` and ``_kernel.py`` and are
anchored by known-answer identities instead (Kernel SHAP of a linear model under
the identity link equals linear SHAP; brute-force Shapley values on 2-4 feature
toys; additivity) - see ``tests/test_oracle_kernel_shap.py``.  The MODEL FUNCTION inside the
Kernel SHAP vectors is pinned: ``tests/golden/kernel_shap_golden.npz``
(``tools/make_golden_kshap.py``) holds coalition expectations computed by the reference's own
unpickled ``LogisticRegression.predict_proba``; only the shap algorithm around it is restated.
"""

from . import scoring, linear_shap, kernel_shap, synth  # noqa: F401
