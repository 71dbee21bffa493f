"""``shap.TreeExplainer`` for the one use the reference makes of it (src/api/cobalt_fast_api.py:46, :100-105):
``explainer = shap.TreeExplainer(xgb_model)``; ``explainer.shap_values(row)[0]``; ``explainer.expected_value``.
Path-dependent TreeSHAP on the margin (shap's default for tree models without background data), computed by the
CUDA kernel K10 (csrc/shap.cuh)."""
from __future__ import annotations

import numpy as np
import torch


class TreeExplainer:
    def __init__(self, model, device: str | None = None):
        self.model = model
        self.booster = model.get_booster() if hasattr(model, "get_booster") else model
        self.device = torch.device(device or getattr(model, "device", None) or "cuda")
        self._expected = None

    def _frame_to_array(self, X) -> np.ndarray:
        if hasattr(X, "columns"):                          # DataFrame: same column checks as predict_proba
            names = self.booster.feature_names
            if names is not None and list(map(str, X.columns)) != list(map(str, names)):
                raise ValueError("feature_names mismatch: " + str(list(X.columns)) + " vs " + str(list(names)))
            X = X.to_numpy(dtype=np.float32, na_value=np.nan) if hasattr(X, "to_numpy") else np.asarray(X)
        X = np.asarray(X, np.float32)
        return X.reshape(1, -1) if X.ndim == 1 else np.ascontiguousarray(X)

    def _contribs(self, X) -> np.ndarray:
        Xn = self._frame_to_array(X)
        with torch.cuda.device(self.device):
            return self.booster.predict_contribs_device(torch.from_numpy(Xn).to(self.device)).cpu().numpy()

    def shap_values(self, X) -> np.ndarray:
        """f32 [n, n_features]: attribution of every feature to the margin (log-odds) of every row."""
        return self._contribs(X)[:, :-1]

    @property
    def expected_value(self) -> float:
        """The margin a row with no known feature gets: cover-weighted tree means + the base margin."""
        if self._expected is None:
            z = np.zeros((1, self.booster.num_features()), np.float32)
            self._expected = float(self._contribs(z)[0, -1])
        return self._expected
