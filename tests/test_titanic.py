"""Config 1 of BASELINE.json: Titanic PassengerDataAll.csv through transmogrify() + sanityCheck(removeBadFeatures),
features as core/src/test/scala/com/salesforce/op/OpWorkflowCVTest.scala:69-82.  tests/golden/titanic.json holds the
raw columns and the oracle's frozen result (made by tests/golden/make_titanic_fixture.py from the reference's file)."""
import json
import math
import os

import numpy as np
import pytest

from oracle import sanity as S
from oracle import vectorizers as V

FIX = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "titanic.json")


def _load():
    with open(FIX) as f:
        return json.load(f)


def _nan(a):
    return np.array([np.nan if v is None else v for v in a], dtype=np.float64)


def test_oracle_reproduces_frozen_titanic_fixture():
    d = _load()
    cols, o = d["columns"], d["oracle"]
    assert d["rows"] == 891 and sum(v is None for v in cols["age"]) == 177 and sum(v is None for v in cols["cabin"]) == 687
    pn, rn = ["sex", "pClass", "cabin"], ["age", "fare"]
    tops = V.pivot_fit([cols[n] for n in pn], 20, 10, True)
    assert tops == o["top_values"] == [["Male", "Female"], ["3", "1", "2"], []]
    xp = V.pivot_transform([cols[n] for n in pn], tops, True, True)
    fills = V.real_vectorizer_fit([cols[n] for n in rn])
    xr = V.real_vectorizer_transform([cols[n] for n in rn], fills, True)
    x, metas = V.combine([xp, xr], [V.pivot_metadata(pn, "PickList", tops, True), V.numeric_metadata(rn, "Real", True)])
    assert [m.col_name() for m in metas] == o["names"] and len(metas) == 15
    res = S.sanity_fit(cols["survived"], x, metas, S.SanityParams(remove_bad_features=True), label_name="survived")
    assert res.dropped == o["dropped"] and res.indices_to_keep == o["indices_to_keep"]
    np.testing.assert_allclose(res.corrs_with_label, _nan(o["corr_label"]), rtol=1e-12, equal_nan=True)
    # README.md:99-110 of the reference reports these on a random 90 % split; on the full file:
    assert abs(res.corrs_with_label[1] - 0.5434) < 1e-4


@pytest.mark.gpu
@pytest.mark.parametrize("device", [None, "cuda"])
def test_titanic_gpu_matches_frozen_oracle_result(device):
    pytest.importorskip("torch")
    from transmogrifai_b200 import features as F
    from transmogrifai_b200.columnar import ColumnarBatch
    from transmogrifai_b200.dsl import fit_and_transform, transmogrify
    d = _load()
    cols, o = d["columns"], d["oracle"]
    data = {"age": (F.REAL, cols["age"]), "sex": (F.PICKLIST, cols["sex"]), "fare": (F.REAL, cols["fare"]),
            "pClass": (F.PICKLIST, cols["pClass"]), "cabin": (F.PICKLIST, cols["cabin"]),
            "survived": (F.REALNN, cols["survived"])}
    batch = ColumnarBatch.from_columns(data, response="survived")
    batch = batch.to(device) if device else batch
    FB = F.FeatureBuilder
    age, sex, fare = FB.Real("age").asPredictor(), FB.PickList("sex").asPredictor(), FB.Real("fare").asPredictor()
    p_class, cabin = FB.PickList("pClass").asPredictor(), FB.PickList("cabin").asPredictor()
    survived = FB.RealNN("survived").asResponse()
    fv = transmogrify([age, sex, fare, p_class, cabin])                 # OpWorkflowCVTest.scala:81
    checked = survived.sanityCheck(fv, removeBadFeatures=True)          # :82, defaults incl. CorrelationLevel.Computed
    out, fitted = fit_and_transform(batch, checked)
    est = checked.origin_stage
    assert est.summary.names[:-1] == o["names"]
    assert est.summary.dropped == o["dropped"]
    assert fitted[-1].indicesToKeep == o["indices_to_keep"]
    pivot_model = [m for m in fitted if hasattr(m, "topValues")][0]
    assert pivot_model.topValues == o["top_values"]
    real_model = [m for m in fitted if hasattr(m, "fillValues")][0]
    np.testing.assert_allclose(real_model.fillValues, o["fills"], rtol=1e-12)
    np.testing.assert_allclose(est.summary.mean, _nan(o["mean"]), rtol=1e-12, atol=1e-300)
    np.testing.assert_allclose(est.summary.variance, _nan(o["variance"]), rtol=1e-9, atol=1e-18)
    np.testing.assert_allclose(est.summary.correlationsWithLabel, _nan(o["corr_label"]), rtol=1e-9, atol=1e-12,
                               equal_nan=True)
    got_v = {g.group: g.cramersV for g in est.summary.categoricalStats}
    assert got_v.keys() == o["cramers_v"].keys()
    for k, v in o["cramers_v"].items():
        assert (v is None and math.isnan(got_v[k])) or abs(got_v[k] - v) <= 1e-9 * abs(v)
    for g in est.summary.categoricalStats:
        assert g.contingencyMatrix.tolist() == o["contingency"][g.group]
    assert out.vectors[checked.name].to_numpy().shape == (891, len(o["indices_to_keep"]))
