"""A/B of the text passes of one config-5 chunk: the span kernel (default), the lane-per-cell kernel
(TMOG_TEXT_PATH=cells) and the row-table kernel (TMOG_TEXT_PATH=rows)."""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
import benchdata  # noqa: E402
from transmogrifai_b200.dsl import fit_and_transform  # noqa: E402

rows = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
batch, feats, label, info = benchdata.make_workload("c5", rows, 1)
checked = bench.build_plan(feats, label, info)
out, fitted = fit_and_transform(batch, checked)
for mode in os.environ.get("AB_MODES", "span,cells,rows").split(","):
    os.environ.pop("TMOG_TEXT_PATH", None)
    if mode != "span":
        os.environ["TMOG_TEXT_PATH"] = mode
    kt = bench.kernel_times("c5", batch, fitted, info, reps=4, text_offset=bench._text_offset(fitted))
    print(mode, [(k["kernel"][:44], round(k["ms"], 2)) for k in kt if "text" in k["kernel"]], flush=True)
