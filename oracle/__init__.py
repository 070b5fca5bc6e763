"""CPU oracle for the transmogrify() + SanityChecker hot path.

TEST INFRASTRUCTURE ONLY.  This package is a plain numpy / pure-Python (plus a
small C helper under ``oracle/c``) restatement of the reference's algorithms,
each function citing the reference file:line it follows.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import it.  The product package
(``transmogrifai_b200``) never imports it and has no CPU fallback.

Pinning status ("how do we know the oracle is right"):

* The reference is 100 % JVM (Scala 2.11 / Spark 2.4.5) and this environment has
  no JVM, so the reference itself cannot be executed here (``oracle/_ref`` does
  not exist for this project).
* Every golden vector / known-answer the reference's own test-suite holds for
  this path is transcribed into ``tests/test_oracle_golden.py`` with its
  file:line, and the oracle must reproduce it (hash indices 107/224, pivot
  vocab order A,C,B / X,Y,Z, cleanString outputs, tokenizer outputs, Cramer's V
  0.2921 / 0.279 / 0.1815, SanityChecker dropped-feature sets, ...).
* Pieces with NO numeric golden in the reference (Pearson correlation values,
  the transmogrified Titanic matrix) are "parity unpinned" beyond definition:
  they are checked against scipy/numpy definitions and frozen as our own
  fixtures under ``tests/golden``.

Third-party algorithms restated here because their source is not vendored in
``/root/reference``: Spark 2.4.5 ``mllib.feature.HashingTF`` (legacy murmur3),
``mllib.stat.Statistics.{colStats,corr,chiSqTest}``, Lucene 7.3.0
``StandardAnalyzer`` (ASCII subset of UAX#29 + Snowball english stop set),
Algebird 0.13.4 map monoid.
"""
