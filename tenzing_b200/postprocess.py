"""Design rules from search results (reference: postprocess/postprocess.py:25-774).

Input: the results CSV the solvers write (``idx|pct01|pct10|pct50|pct90|pct99|stddev|<op
json>|...``, optional JSON header lines).  Steps, as in the reference:

1. sort implementations by pct10; convolve the sorted times with a step kernel and take the
   prominent peaks as boundaries between *performance classes*;
2. turn every sequence into boolean features "a before b" and "a same stream as b" over the
   program's (non-synchronisation) ops, drop constant and duplicate columns;
3. fit decision trees (entropy, balanced classes) with a growing leaf budget until the training
   accuracy stops improving, and print the root-to-leaf paths as rules;
4. ``evaluate_rules``: how well do rules learned from the first n search iterations classify
   all implementations (does the search find the structure early?).

matplotlib is optional (no plots without it); pandas / scipy / scikit-learn are required.
"""
import json
import math
import sys

import numpy as np

SYNC_KINDS = {"CudaEventRecord", "CudaEventSync", "CudaStreamWaitEvent", "StreamWait", "StreamSync"}


def read_results(path_or_text):
    """-> (times[n,6] as pct01,pct10,pct50,pct90,pct99,stddev ; sequences: list of list of op dicts)"""
    text = open(path_or_text).read() if "\n" not in path_or_text else path_or_text
    times, seqs = [], []
    for line in text.splitlines():
        if not line or line[0] in "{#":
            continue
        cols = line.split("|")
        if len(cols) < 8:
            continue
        try:
            t = [float(c) for c in cols[1:7]]
        except ValueError:
            continue  # a textual header row
        times.append(t)
        seqs.append([json.loads(c) for c in cols[7:] if c])
    return np.array(times, dtype=np.float64).reshape(-1, 6), seqs


def find_classes(pct10, radius_fraction=0.005, prominence_pct=99.0):
    """boundaries (indices into the sorted order) between performance classes, and labels"""
    from scipy.signal import find_peaks
    order = np.argsort(pct10, kind="stable")
    arr = np.asarray(pct10)[order]
    kr = max(1, int(math.ceil(len(arr) * radius_fraction)))
    if len(arr) < 2 * kr + 1:
        return order, np.array([], dtype=int), np.zeros(len(arr), dtype=int)
    kernel = np.array([1.0] * kr + [-1.0] * kr)
    res = np.convolve(arr, kernel, "valid")
    cutoff = np.percentile(res, prominence_pct)
    peaks, _ = find_peaks(res, prominence=max(cutoff, 1e-300), width=1)
    peaks = peaks + len(kernel) // 2
    labels_sorted = np.searchsorted(peaks, np.arange(len(arr)), side="right")
    labels = np.empty(len(arr), dtype=int)
    labels[order] = labels_sorted
    return order, peaks, labels


def program_ops(seq):
    """names of the non-synchronisation ops of a sequence, with their streams"""
    return [(o["name"], o.get("stream")) for o in seq if o.get("kind") not in SYNC_KINDS]


def featurize(seqs):
    """boolean feature matrix: 'a before b' for every unordered pair, 'a same stream as b' for
    every pair of GPU ops; ops missing from a sequence give False"""
    names = sorted({n for s in seqs for n, _ in program_ops(s)} - {"Start", "Finish"})
    gpu = sorted({n for s in seqs for n, st in program_ops(s) if st is not None})
    feats, cols = [], []
    pos = [{n: i for i, (n, _) in enumerate(program_ops(s))} for s in seqs]
    stream = [{n: st for n, st in program_ops(s)} for s in seqs]
    for i, a in enumerate(names):
        for b in names[i + 1:]:
            cols.append(f"{a} before {b}")
            feats.append([p.get(a, 1 << 30) < p.get(b, 1 << 30) for p in pos])
    for i, a in enumerate(gpu):
        for b in gpu[i + 1:]:
            cols.append(f"{a} same stream as {b}")
            feats.append([(a in s and b in s and s[a] == s[b]) for s in stream])
    X = np.array(feats, dtype=bool).T if feats else np.zeros((len(seqs), 0), dtype=bool)
    return X, cols


def prune_features(X, cols):
    """drop constant columns and exact duplicates (the first of a duplicate group is kept)"""
    keep, seen = [], set()
    for j in range(X.shape[1]):
        col = X[:, j]
        if col.all() or not col.any():
            continue
        key = col.tobytes()
        inv = (~col).tobytes()  # "a before b" and its negation carry the same information
        if key in seen or inv in seen:
            continue
        seen.add(key)
        keep.append(j)
    return X[:, keep], [cols[j] for j in keep]


def fit_tree(X, y, max_leaf_nodes=None, seed=0):
    from sklearn.tree import DecisionTreeClassifier
    clf = DecisionTreeClassifier(criterion="entropy", class_weight="balanced",
                                 max_leaf_nodes=max_leaf_nodes, random_state=seed)
    return clf.fit(X, y)


def grow_tree(X, y, seed=0):
    """smallest leaf budget after which the training accuracy no longer improves"""
    n_classes = len(set(y))
    best, best_acc = None, -1.0
    for leaves in range(max(2, n_classes), max(2, n_classes) + 32):
        clf = fit_tree(X, y, leaves, seed)
        acc = clf.score(X, y)
        if acc > best_acc + 1e-12:
            best, best_acc = clf, acc
        elif best is not None:
            break
        if acc >= 1.0:
            break
    return best, best_acc


def tree_rules(clf, cols):
    """one rule per leaf: (conditions, class, samples)"""
    t = clf.tree_
    rules = []

    def walk(node, conds):
        if t.children_left[node] == t.children_right[node]:
            counts = t.value[node][0]
            rules.append((list(conds), int(clf.classes_[int(np.argmax(counts))]), int(t.n_node_samples[node])))
            return
        name = cols[t.feature[node]]
        walk(t.children_left[node], conds + [f"NOT ({name})"])   # feature <= 0.5  -> False
        walk(t.children_right[node], conds + [name])
    walk(0, [])
    return rules


def rules_text(rules, class_times=None):
    out = []
    for conds, cls, n in sorted(rules, key=lambda r: r[1]):
        head = f"class {cls}"
        if class_times is not None and cls in class_times:
            lo, hi = class_times[cls]
            head += f" [{lo:.3e} s .. {hi:.3e} s]"
        out.append(f"{head} ({n} implementations): " + (" AND ".join(conds) if conds else "always"))
    return "\n".join(out)


def analyse(path_or_text, prominence_pct=99.0, seed=0):
    """full pipeline -> dict(labels, peaks, features, tree, rules, accuracy, text)"""
    times, seqs = read_results(path_or_text)
    if len(seqs) == 0:
        raise ValueError("no result rows")
    pct10 = times[:, 1]
    order, peaks, labels = find_classes(pct10, prominence_pct=prominence_pct)
    X, cols = featurize(seqs)
    X, cols = prune_features(X, cols)
    class_times = {int(c): (float(pct10[labels == c].min()), float(pct10[labels == c].max()))
                   for c in set(labels)}
    if X.shape[1] == 0 or len(set(labels)) < 2:
        return dict(labels=labels, peaks=peaks, features=cols, tree=None, rules=[], accuracy=1.0,
                    class_times=class_times, text="(a single performance class or no varying feature)")
    clf, acc = grow_tree(X, labels, seed)
    rules = tree_rules(clf, cols)
    return dict(labels=labels, peaks=peaks, features=cols, tree=clf, rules=rules, accuracy=acc,
                class_times=class_times, text=rules_text(rules, class_times), X=X)


def evaluate_rules(path_or_text, prefixes=(50, 100, 200, 400), prominence_pct=99.0, seed=0):
    """accuracy on ALL implementations of trees trained on the first n rows (search order)"""
    times, seqs = read_results(path_or_text)
    pct10 = times[:, 1]
    _, _, labels = find_classes(pct10, prominence_pct=prominence_pct)
    X, cols = featurize(seqs)
    X, cols = prune_features(X, cols)
    out = {}
    for n in list(prefixes) + [len(seqs)]:
        n = min(n, len(seqs))
        if n < 2 or len(set(labels[:n])) < 2 or X.shape[1] == 0:
            continue
        clf, _ = grow_tree(X[:n], labels[:n], seed)
        out[n] = float(clf.score(X, labels))
    return out


def classes_svg(pct10, peaks, width=900, height=360):
    """The picture the reference draws with matplotlib (sorted times, a vertical line at every
    class boundary; postprocess.py:254-316), as a dependency-free SVG string: implementations
    sorted by pct10 on the x axis, time on a log y axis, class boundaries as dashed lines."""
    arr = np.sort(np.asarray(pct10, dtype=np.float64))
    n = len(arr)
    ml, mr, mt, mb = 70, 15, 20, 40
    w, h = width - ml - mr, height - mt - mb
    lo, hi = max(arr.min(), 1e-300), max(arr.max(), 1e-300)
    if hi <= lo:
        hi = lo * 1.0001
    llo, lhi = math.log10(lo), math.log10(hi)

    def px(i):
        return ml + (w * i / max(n - 1, 1))

    def py(t):
        return mt + h * (1.0 - (math.log10(max(t, 1e-300)) - llo) / (lhi - llo))

    out = [f'<svg xmlns="http://www.w3.org/2000/svg" width="{width}" height="{height}" '
           f'viewBox="0 0 {width} {height}" font-family="sans-serif" font-size="11">',
           f'<rect x="{ml}" y="{mt}" width="{w}" height="{h}" fill="white" stroke="#444"/>']
    for k in range(6):                                       # y ticks, evenly spaced in log space
        t = 10 ** (llo + (lhi - llo) * k / 5)
        y = py(t)
        out.append(f'<line x1="{ml - 4}" y1="{y:.1f}" x2="{ml}" y2="{y:.1f}" stroke="#444"/>')
        out.append(f'<text x="{ml - 6}" y="{y + 4:.1f}" text-anchor="end">{t * 1e6:.1f} us</text>')
    for p_ in peaks:                                         # class boundaries
        x = px(int(p_))
        out.append(f'<line x1="{x:.1f}" y1="{mt}" x2="{x:.1f}" y2="{mt + h}" stroke="#c33" '
                   'stroke-dasharray="4 3"/>')
    pts = " ".join(f"{px(i):.1f},{py(t):.1f}" for i, t in enumerate(arr))
    out.append(f'<polyline points="{pts}" fill="none" stroke="#1a5fb4" stroke-width="1.5"/>')
    out.append(f'<text x="{ml + w / 2:.0f}" y="{height - 10}" text-anchor="middle">'
               f'{n} implementations sorted by pct10; {len(peaks) + 1} performance classes</text>')
    out.append("</svg>")
    return "\n".join(out)


def export_dot(clf, cols, path):
    from sklearn.tree import export_graphviz
    export_graphviz(clf, out_file=path, feature_names=cols, filled=True, rounded=True)


def main(argv=None):
    argv = sys.argv[1:] if argv is None else argv
    if not argv:
        print("usage: python -m tenzing_b200.postprocess results.csv [out_prefix]")
        return 2
    res = analyse(argv[0])
    prefix = argv[1] if len(argv) > 1 else argv[0] + "."
    print(f"{len(res['labels'])} implementations, {len(set(res['labels']))} performance classes, "
          f"{len(res['features'])} features, tree accuracy {res['accuracy']:.3f}")
    print(res["text"])
    with open(prefix + "rules.txt", "w") as f:
        f.write(res["text"] + "\n")
    if res["tree"] is not None:
        export_dot(res["tree"], res["features"], prefix + "tree.dot")
    times, _ = read_results(argv[0])
    with open(prefix + "classes.svg", "w") as f:
        f.write(classes_svg(times[:, 1], res["peaks"]))
    ev = evaluate_rules(argv[0])
    if ev:
        print("rules from the first n iterations -> accuracy on all:", ev)
    return 0


if __name__ == "__main__":
    sys.exit(main())
