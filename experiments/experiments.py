"""Script form of the reference's ``experiments.ipynb`` (cells 1, 4, 5-7) for this framework.

* cell 1  (``experiments.ipynb:58-70``): 4-rank parallel tree on iris[:, :2], depth 3, printed by rank 0
* cell 4  (``:184-189``): serial depth-5 tree, ``precision=1``
* cell 5  (``:198-209``) + ``time_data.csv``: fit wall-time for n = 1, 11, ..., 241 distinct-class rows,
  serial and with k = 2, 5, 8 ranks; written to ``experiments/time_data.csv`` (rows: serial, k=2, k=5, k=8)
* cell 3  (``:99-137``): decision boundaries for depth 2 and 5, written as PNG files by a tiny built-in writer
* cell 7 (timing plot) needs matplotlib; it is drawn only when that is importable.

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


def _write_png(path, rgb):
    """Minimal PNG writer (zlib + struct) so the figures need no plotting library."""
    import struct
    import zlib

    h, w, _ = rgb.shape
    raw = b"".join(b"\x00" + rgb[i].astype(np.uint8).tobytes() for i in range(h))

    def chunk(tag, data):
        c = struct.pack(">I", len(data)) + tag + data
        return c + struct.pack(">I", zlib.crc32(tag + data) & 0xFFFFFFFF)

    png = b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, 2, 0, 0, 0))
    png += chunk(b"IDAT", zlib.compress(raw, 9)) + chunk(b"IEND", b"")
    pathlib.Path(path).write_bytes(png)


def cell3(res: int = 300):
    """Decision boundaries for max_depth 2 and 5 on iris[:, :2] (``experiments.ipynb:99-137``), as PNG files."""
    iris = load_iris()
    X, y = iris.data[:, :2], iris.target
    fill = np.array([[0x97, 0xC4, 0x77], [0xFD, 0x91, 0x77], [0x97, 0x91, 0xDD]])
    dots = np.array([[0x00, 0x80, 0x00], [0xFF, 0x00, 0x00], [0x00, 0x00, 0xFF]])
    x0, x1 = X[:, 0].min() - 0.5, X[:, 0].max() + 0.5
    y0, y1 = X[:, 1].min() - 0.5, X[:, 1].max() + 0.5
    gx, gy = np.meshgrid(np.linspace(x0, x1, res), np.linspace(y1, y0, res))
    for depth in (2, 5):
        clf = DecisionTreeClassifier(max_depth=depth, device="cpu").fit(X, y)
        img = fill[clf.predict(np.c_[gx.ravel(), gy.ravel()])].reshape(res, res, 3).copy()
        px = ((X[:, 0] - x0) / (x1 - x0) * (res - 1)).astype(int)
        py = ((y1 - X[:, 1]) / (y1 - y0) * (res - 1)).astype(int)
        for dx in (-1, 0, 1):
            for dy in (-1, 0, 1):
                img[np.clip(py + dy, 0, res - 1), np.clip(px + dx, 0, res - 1)] = dots[y]
        _write_png(HERE / f"decision_boundary_depth{depth}.png", img)
        print(f"wrote decision_boundary_depth{depth}.png (training accuracy {clf.score(X, y):.4f})")


def cell5(reps: int = 3):
    """Timing table.  Parallel rows come from ``parallel_timing.py`` (real processes under torchrun)
    when its CSV files are present; thread-ranks share one interpreter lock, so timing them would
    measure the GIL rather than the algorithm and is only the fallback."""
    x_dim = np.arange(1, 250, 10)
    rows = []
    for k in (1, 2, 5, 8):
        measured = HERE / f"time_data_k{k}_processes.csv"
        if k > 1 and measured.exists():
            out = [float(v) for v in measured.read_text().split(",")]
            rows.append(out)
            print(f"k={k} ({measured.name}):", out)
            continue
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


def _plot_without_matplotlib(x_dim, rows, w: int = 640, h: int = 420):
    """Line chart of the four timing series drawn straight into an RGB array (no plotting library)."""
    img = np.full((h, w, 3), 255, dtype=np.uint8)
    left, right, top, bottom = 50, 20, 20, 40
    colors = [(0x1F, 0x77, 0xB4), (0xFF, 0x7F, 0x0E), (0x2C, 0xA0, 0x2C), (0xD6, 0x27, 0x28)]
    xs = np.asarray(x_dim, dtype=float)
    ymax = max(max(r) for r in rows) * 1.05 or 1.0

    def to_px(x, y):
        px = left + (x - xs[0]) / max(xs[-1] - xs[0], 1e-9) * (w - left - right)
        py = h - bottom - y / ymax * (h - top - bottom)
        return px, py

    img[h - bottom, left : w - right] = 0  # axes
    img[top : h - bottom, left] = 0
    for frac in (0.25, 0.5, 0.75, 1.0):  # dashed horizontal grid
        gy = int(h - bottom - frac * (h - top - bottom))
        img[gy, left : w - right : 4] = 200
    for r, col in zip(rows, colors):
        pts = [to_px(x, y) for x, y in zip(xs, r)]
        for (ax_, ay_), (bx_, by_) in zip(pts, pts[1:]):
            steps = int(max(abs(bx_ - ax_), abs(by_ - ay_))) + 1
            lx = np.linspace(ax_, bx_, steps).round().astype(int)
            ly = np.linspace(ay_, by_, steps).round().astype(int)
            img[np.clip(ly, 0, h - 1), np.clip(lx, 0, w - 1)] = col
        for px, py in pts:  # markers
            img[int(py) - 2 : int(py) + 3, int(px) - 2 : int(px) + 3] = col
    _write_png(HERE / "timing.png", img)
    print(f"wrote timing.png (series: serial, k=2, k=5, k=8; x = n_samples {int(xs[0])}..{int(xs[-1])}, "
          f"y = 0..{ymax:.1f} ms)")


def plots(x_dim, rows):
    try:
        import matplotlib

        matplotlib.use("Agg")
        import matplotlib.pyplot as plt
    except ImportError:
        _plot_without_matplotlib(x_dim, rows)
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
    cell3()
    print()
    plots(*cell5())
