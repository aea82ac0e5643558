"""Script form of the reference's ``experiments.ipynb`` (cells 1, 4, 5-7) for this framework.

* cell 1  (``experiments.ipynb:58-70``): 4-rank parallel tree on iris[:, :2], depth 3, printed by rank 0
* cell 4  (``:184-189``): serial depth-5 tree, ``precision=1``
* cell 5  (``:198-209``) + ``time_data.csv``: fit wall-time for n = 1, 11, ..., 241 distinct-class rows,
  serial and with k = 2, 5, 8 ranks; written to ``experiments/time_data.csv`` (rows: serial, k=2, k=5, k=8)
* cells 3, 7 (plots) need matplotlib; they are drawn only when it is importable.

Ranks are threads of this process (``ThreadComm``) so the script runs anywhere; under
``torchrun`` the same classes use the real process group.
"""
import pathlib
import sys
import time

import numpy as np
from sklearn.datasets import load_iris

sys.path.insert(0, str(pathlib.Path(__file__).resolve().parents[1]))
from mpitree_b200.parallel.comm import run_threaded  # noqa: E402
from mpitree_b200.tree import DecisionTreeClassifier, ParallelDecisionTreeClassifier  # noqa: E402

HERE = pathlib.Path(__file__).resolve().parent


def cell1():
    iris = load_iris()
    X, y = iris.data[:, :2], iris.target

    def rank_main(comm):
        clf = ParallelDecisionTreeClassifier(max_depth=3, comm=comm, device="cpu").fit(X, y)
        if not comm.rank:
            print(clf.export_text(feature_names=iris.feature_names, class_names=iris.target_names))

    run_threaded(4, rank_main)


def cell4():
    iris = load_iris()
    X, y = iris.data[:, :2], iris.target
    clf = DecisionTreeClassifier(max_depth=5, device="cpu").fit(X, y)
    print(clf.export_text(feature_names=iris.feature_names, class_names=iris.target_names, precision=1))


def cell5(reps: int = 3):
    x_dim = np.arange(1, 250, 10)
    rows = []
    for k in (1, 2, 5, 8):
        out = []
        for n in x_dim:
            X = np.arange(n).reshape(-1, 1)
            y = np.arange(n)
            best = float("inf")
            for _ in range(reps):
                t0 = time.perf_counter()
                if k == 1:
                    DecisionTreeClassifier(device="cpu").fit(X, y)
                else:
                    run_threaded(k, lambda comm: ParallelDecisionTreeClassifier(comm=comm, device="cpu").fit(X, y))
                best = min(best, (time.perf_counter() - t0) * 1000)
            out.append(round(best, 2))
        rows.append(out)
        print(f"k={k}:", out)
    (HERE / "time_data.csv").write_text("\n".join(", ".join(str(v) for v in r) for r in rows) + "\n")
    return x_dim, rows


def plots(x_dim, rows):
    try:
        import matplotlib

        matplotlib.use("Agg")
        import matplotlib.pyplot as plt
    except ImportError:
        print("matplotlib not installed: skipping the figures")
        return
    fig = plt.figure()
    ax = fig.add_subplot()
    for r, label in zip(rows, ("serial", "k=2", "k=5", "k=8")):
        ax.plot(x_dim, r, marker="o", label=label)
    ax.set_xlabel("n_samples")
    ax.set_ylabel("time (ms)")
    plt.title("Execution Times")
    plt.legend(loc=2)
    plt.grid(ls="--")
    fig.savefig(HERE / "timing.png")


if __name__ == "__main__":
    cell1()
    print()
    cell4()
    print()
    plots(*cell5())
