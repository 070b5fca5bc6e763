"""Generates tests/golden/titanic.json from the reference's own test data (run in the build container only; the GPU
box has no /root/reference).

Input : /root/reference/test-data/PassengerDataAll.csv (891 rows; columns id, survived, pClass, name, sex, age, sibSp,
        parCh, ticket, fare, cabin, embarked), features as core/src/test/scala/com/salesforce/op/OpWorkflowCVTest.scala:69-82
        (age, fare: Real; sex, pClass, cabin: PickList; label survived.occurs(): RealNN).
Output: the six raw columns plus the ORACLE's result for Seq(age, sex, fare, pClass, cabin).transmogrify() followed by
        sanityCheck(removeBadFeatures = true) with the reference defaults (checkSample 1.0; 891 rows < sampleLowerLimit
        1000, so fraction() == 1.0 and nothing is sampled).  The reference asserts only model quality on this data, so
        this fixture is OUR frozen golden (parity unpinned for the Titanic matrix itself).
"""
import csv
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import sanity as S  # noqa: E402
from oracle import vectorizers as V  # noqa: E402

SRC = "/root/reference/test-data/PassengerDataAll.csv"


def main():
    rows = list(csv.reader(open(SRC, newline="")))
    opt = lambda s: None if s == "" else s  # noqa: E731
    num = lambda s: None if s == "" else float(s)  # noqa: E731
    cols = {
        "survived": [float(r[1]) for r in rows],
        "pClass": [opt(r[2]) for r in rows],
        "sex": [opt(r[4]) for r in rows],
        "age": [num(r[5]) for r in rows],
        "fare": [num(r[9]) for r in rows],
        "cabin": [opt(r[10]) for r in rows],
    }
    out = {"source": SRC, "rows": len(rows), "columns": cols, "oracle": oracle_result(cols)}
    with open(os.path.join(ROOT, "tests", "golden", "titanic.json"), "w") as f:
        json.dump(out, f)
    print("rows", len(rows), "dropped", out["oracle"]["dropped"])


def oracle_result(cols):
    """Type groups in type-name order (PickList < Real), members in input order (Transmogrifier.scala:114)."""
    pnames, rnames = ["sex", "pClass", "cabin"], ["age", "fare"]
    pcols, rcols = [cols[n] for n in pnames], [cols[n] for n in rnames]
    tops = V.pivot_fit(pcols, 20, 10, True)
    xp = V.pivot_transform(pcols, tops, True, True)
    fills = V.real_vectorizer_fit(rcols)
    xr = V.real_vectorizer_transform(rcols, fills, True)
    x, metas = V.combine([xp, xr], [V.pivot_metadata(pnames, "PickList", tops, True),
                                    V.numeric_metadata(rnames, "Real", True)])
    p = S.SanityParams(remove_bad_features=True, feature_feature_corr_level="computed")
    res = S.sanity_fit(cols["survived"], x, metas, p, label_name="survived")
    nan = lambda a: [None if v != v else float(v) for v in a]  # noqa: E731
    return {"top_values": tops, "fills": fills, "names": [m.col_name() for m in metas], "dropped": res.dropped,
            "indices_to_keep": res.indices_to_keep, "mean": nan(res.col_stats.mean),
            "variance": nan(res.col_stats.variance), "corr_label": nan(res.corrs_with_label),
            "cramers_v": {g.group: (None if g.cramers_v != g.cramers_v else g.cramers_v) for g in res.cat_stats},
            "contingency": {g.group: g.contingency_matrix.tolist() for g in res.cat_stats}}


if __name__ == "__main__":
    main()
