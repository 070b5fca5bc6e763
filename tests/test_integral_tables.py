"""Host logic of the chunk-streamed IntegralVectorizer fit: value -> count maps of chunks / ranks add up key-wise and the
mode is the most frequent value, ties to the smaller one (SequenceAggregators.scala:100-131, ModeSeqNullInt)."""
import numpy as np

from oracle import vectorizers as V


def test_value_count_maps_merge_like_the_oracle():
    from transmogrifai_b200.stages.integral import INT64_MIN, _add_tables, _merge_tables
    rng = np.random.default_rng(4)
    cols = [[None if rng.random() < 0.2 else int(v) for v in rng.choice([5, -2, 9, INT64_MIN, 2 ** 40], size=600)],
            [None] * 600,
            [int(v) for v in rng.integers(-3, 4, size=600)]]
    bounds = [0, 1, 250, 251, 600]

    def table(col):
        vals = np.array([v for v in col if v is not None], dtype=np.int64)
        k, w = np.unique(vals, return_counts=True) if len(vals) else (np.zeros(0, np.int64), np.zeros(0, np.int64))
        return k, w.astype(np.int64)

    acc = None
    per_chunk = []
    for a, b in zip(bounds[:-1], bounds[1:]):
        local = [table(c[a:b]) for c in cols]
        per_chunk.append(local)
        acc = _add_tables(acc, local)
    exp = V.mode_seq_null_int(cols)
    assert _merge_tables([acc]) == exp                       # chunks added one after the other
    assert _merge_tables(per_chunk) == exp                   # sources (ranks) merged at once
    assert exp[1] == 0                                       # no value at all -> 0
    # a tie goes to the smaller value
    tie = [(np.array([7, 3], dtype=np.int64), np.array([4, 4], dtype=np.int64))]
    assert _merge_tables([tie]) == [3]
