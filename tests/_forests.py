"""Random forests in the wire format for the scoring / TreeSHAP tests (CPU and GPU suites)."""
import numpy as np


def random_forest(rng, n_trees, n_features, max_depth, n_thr):
    """Random trees in wire format (ragged depth), thresholds from a small per-feature pool so codes collide."""
    pools = [np.sort(rng.normal(size=rng.integers(1, n_thr + 1)).astype(np.float32)) for _ in range(n_features)]
    left, right, feat, cond, dl, offs = [], [], [], [], [], [0]
    for _ in range(n_trees):
        nodes = []                                   # [left, right, feat, cond, dl]

        def grow(d):
            i = len(nodes)
            nodes.append(None)
            if d == max_depth or (d > 0 and rng.random() < 0.2):
                nodes[i] = [-1, -1, 0, np.float32(rng.normal() * 0.1), 0]
                return i
            j = int(rng.integers(n_features))
            rec = [0, 0, j, pools[j][rng.integers(len(pools[j]))], int(rng.random() < 0.5)]
            nodes[i] = rec
            rec[0] = grow(d + 1)
            rec[1] = grow(d + 1)
            return i
        grow(0)
        for r in nodes:
            left.append(r[0]); right.append(r[1]); feat.append(r[2]); cond.append(r[3]); dl.append(r[4])
        offs.append(len(left))
    f = {"tree_offsets": np.array(offs, np.int64), "left": np.array(left, np.int32), "right": np.array(right, np.int32),
         "split_index": np.array(feat, np.int32), "split_cond": np.array(cond, np.float32),
         "default_left": np.array(dl, np.uint8)}
    n = len(left)
    f["parent"] = np.zeros(n, np.int32)
    for k in ("base_weight", "loss_chg"):
        f[k] = np.zeros(n, np.float32)
    cover = np.zeros(n, np.float64)                  # consistent covers: leaves draw one, parents add their children
    for t in range(n_trees):
        o, e = offs[t], offs[t + 1]
        for i in range(e - 1, o - 1, -1):            # children always follow their parent in this numbering
            if f["left"][i] == -1:
                cover[i] = rng.integers(1, 50) * 0.25
            else:
                cover[i] = cover[o + f["left"][i]] + cover[o + f["right"][i]]
    f["sum_hess"] = cover.astype(np.float32)
    return f, pools
