"""Mirror of the reference's scoring service (src/api/cobalt_fast_api.py) on the B200 path.

Same module-level names (``app``, ``xgb_model``, ``SingleInput``, ``BulkInput``, ``predict_proba_df``), same three
routes and JSON shapes (reference :96-108, :113-126, :128-143).  Differences, all outside the hot path:
the model comes from a local file (``COBALT_MODEL_PATH``) instead of S3 (:36-54), and ``shap.TreeExplainer`` is this
package's ``explain.TreeExplainer`` (path-dependent TreeSHAP on the GPU).
The model file may be a pickle of this package's XGBClassifier or the reference's own xgboost pickle / ``.ubj``
(decoded without xgboost, see ``ubjson.py``).
"""
from __future__ import annotations

import os
from contextlib import asynccontextmanager
from io import BytesIO
from pathlib import Path
from typing import Dict, List

import numpy as np
import pandas as pd
from fastapi import FastAPI, File, HTTPException, UploadFile
from pydantic import BaseModel, ConfigDict, Field

from ..explain import TreeExplainer

local_model_path = Path(os.environ.get("COBALT_MODEL_PATH", "models/xgb_model_tree.pkl"))

app = FastAPI(title="Cobalt XGBoost Inference API")
xgb_model = None
explainer = None


def load_model(path) -> object:
    """An xgboost pickle / .ubj is decoded without xgboost and WITHOUT executing the pickle (``ubjson.py`` scans the
    opcodes for the booster payload); only a file that holds no booster payload -- i.e. this package's own estimator
    pickle -- is handed to ``joblib.load``."""
    import joblib

    from ..ubjson import NoBoosterPayload, load_xgboost_pickle
    try:
        return load_xgboost_pickle(str(path))
    except NoBoosterPayload:
        pass
    model = joblib.load(path)
    if not (hasattr(model, "predict_proba") and type(model).__module__.startswith("cobalt_smart_lender_ai_b200")):
        raise TypeError(f"{path}: neither an xgboost booster payload nor a pickle of this package's XGBClassifier")
    return model


@asynccontextmanager
async def lifespan(app: FastAPI):
    global xgb_model, explainer
    try:
        print(f"[INFO] Loading model from {local_model_path}")
        xgb_model = load_model(local_model_path)
        explainer = TreeExplainer(xgb_model)               # reference :46: shap.TreeExplainer(xgb_model)
        print("[INFO] Model and SHAP Explainer ready.")
    except Exception as e:
        print(f"[ERROR] Model load failed: {e}")
        raise RuntimeError("Failed to load model.")
    yield

app.router.lifespan_context = lifespan


class SingleInput(BaseModel):
    model_config = ConfigDict(populate_by_name=True)

    loan_amnt: float
    term: float
    installment: float
    fico_range_low: float
    last_fico_range_high: float
    open_il_12m: float
    open_il_24m: float
    max_bal_bc: float
    num_rev_accts: float
    pub_rec_bankruptcies: float
    emp_length_num: float
    earliest_cr_line_days: float
    grade_E: int
    home_ownership_MORTGAGE: int
    verification_status_Verified: int
    application_type_Joint_App: int = Field(alias="application_type_Joint App")
    hardship_status_BROKEN: int
    hardship_status_COMPLETE: int
    hardship_status_COMPLETED: int
    hardship_status_No_Hardship: int = Field(alias="hardship_status_No Hardship")


class BulkInput(BaseModel):
    data: List[Dict]


def predict_proba_df(df: pd.DataFrame) -> np.ndarray:
    return xgb_model.predict_proba(df)[:, 1]


@app.post("/predict")
def predict_single(input_data: SingleInput):
    row = pd.DataFrame([input_data.model_dump(by_alias=True)])
    y_proba = predict_proba_df(row)[0]
    shap_vals = explainer.shap_values(row)[0].tolist()
    return {
        "prob_default": float(y_proba),
        "shap_values": shap_vals,
        "base_value": float(explainer.expected_value),
        "features": row.columns.tolist(),
        "input_row": row.iloc[0].to_dict(),
    }


@app.post("/predict_bulk_csv")
async def predict_bulk_csv(file: UploadFile = File(...)):
    try:
        contents = await file.read()
        df = pd.read_csv(BytesIO(contents))
        df["prob_default"] = predict_proba_df(df)
        # reference :121: replace([inf, -inf], nan).fillna("null") -- the mask must come from the REPLACED frame
        d = df.replace([np.inf, -np.inf], np.nan)
        df_clean = d.astype(object).where(d.notna(), "null")
        return {"predictions": df_clean.to_dict(orient="records")}
    except Exception as e:
        print(f"[ERROR] Bulk prediction failed: {e}")
        raise HTTPException(status_code=500, detail=f"Bulk prediction failed: {e}")


@app.post("/feature_importance_bulk")
def feature_importance_bulk(data: BulkInput):
    if not data.data:
        raise HTTPException(status_code=400, detail="No data provided.")
    try:
        booster = xgb_model.get_booster()
        importance_dict = booster.get_score(importance_type="gain")
        sorted_features = sorted(importance_dict.items(), key=lambda x: x[1], reverse=True)
        return {"top_features": [{"feature": k, "importance": v} for k, v in sorted_features[:10]]}
    except Exception as e:
        raise HTTPException(status_code=500, detail=f"Feature importance computation failed: {e}")


if __name__ == "__main__":
    import uvicorn
    uvicorn.run("cobalt_smart_lender_ai_b200.api.cobalt_fast_api:app", host="0.0.0.0", port=8000)
