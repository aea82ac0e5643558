"""Same-machine CPU comparison against the unmodified reference on workloads it can run."""
import sys
import time

import numpy as np

sys.path.insert(0, ".")
from baseline.oracle import load_reference  # noqa: E402
from mpitree_b200.tree import DecisionTreeClassifier  # noqa: E402

ref = load_reference()


def best(fn, reps=3):
    out = []
    for _ in range(reps):
        t0 = time.perf_counter()
        fn()
        out.append(time.perf_counter() - t0)
    return min(out)


print("| workload | reference (s) | mpitree_b200 CPU engine (s) | speed-up | same tree |")
print("|---|---|---|---|---|")
for n in (41, 121, 241):
    X, y = np.arange(n).reshape(-1, 1), np.arange(n)
    a = best(lambda: ref.DecisionTreeClassifier().fit(X, y))
    b = best(lambda: DecisionTreeClassifier(device="cpu").fit(X, y))
    same = ref.DecisionTreeClassifier().fit(X, y).export_text() == DecisionTreeClassifier(device="cpu").fit(X, y).export_text()
    print(f"| notebook cell 5, n={n} distinct classes, 1 feature | {a:.4f} | {b:.4f} | {a / b:.1f}x | {same} |")
rng = np.random.default_rng(0)
for n in (1000, 2000, 4000):
    X = rng.normal(size=(n, 8))
    y = (X[:, 0] + 0.5 * X[:, 3] + 0.3 * rng.normal(size=n) > 0).astype(int)
    reps = 1 if n > 1000 else 2
    a = best(lambda: ref.DecisionTreeClassifier(max_depth=4).fit(X, y), reps)
    b = best(lambda: DecisionTreeClassifier(max_depth=4, max_bins=256, device="cpu").fit(X, y), 3)
    print(f"| continuous {n} x 8, 2 classes, max_depth=4 (ours: 256 quantile bins) | {a:.2f} | {b:.4f} | {a / b:.0f}x | binned |")
