"""Pins oracle/per_oracle.c against vectors produced by the reference's own
njit code (tests/golden/make_golden.py).  Bit-exact for indices, sums, mins."""
import hashlib

import numpy as np
import pytest

from oracle import per_oracle as po


def _sha(a):
    return np.frombuffer(hashlib.sha256(np.ascontiguousarray(a).tobytes()).digest(), np.uint8)


def _replay_tape(g, prefix="", fill_p=None):
    cap, batch = int(g[prefix + "cap"]), int(g[prefix + "batch"])
    tree = po.OracleSumTree(cap)
    fill_idx = g[prefix + "fill_idx"]
    fill_p = g[prefix + "fill_p"] if fill_p is None else fill_p
    for s in range(0, len(fill_idx), 64):
        tree.update(fill_idx[s:s + 64], fill_p[s:s + 64])
    for t in range(len(g[prefix + "u"])):
        idx, val = tree.sample(batch, g[prefix + "u"][t])
        assert np.array_equal(idx, g[prefix + "idx"][t]), f"step {t}"
        assert np.array_equal(val, g[prefix + "val"][t])
        root = g[prefix + "root"][t]
        assert tree.sum() == root[0] and tree.min(0) == root[1]
        tree.update(idx, g[prefix + "newp"][t])
    fr = g[prefix + "final_root"]
    assert tree.sum() == fr[0] and tree.min(0) == fr[1]
    return tree


@pytest.mark.parametrize("name", ["tree_cap1000_full", "tree_cap4096_sparse"])
def test_tape_full_arrays(golden, name):
    g = golden(name)
    tree = _replay_tape(g)
    assert np.array_equal(tree._sum, g["sum"])
    assert np.array_equal(tree._min, g["min"])
    assert np.array_equal(tree._val, g["valarr"])


def test_min_quirk_visible(golden):
    # all priorities > 1 on a sparse tree: reference min is clamped to <= 1
    g = golden("tree_cap4096_sparse")
    assert g["root"][:, 1].max() <= 1.0
    assert g["fill_p"].min() > 1.0


def test_edge_capacities(golden):
    g = golden("tree_edges")
    for cap in g["caps"]:
        pre = f"c{cap}_"
        tree = _replay_tape(g, pre)
        assert np.array_equal(tree._sum, g[pre + "sum"])
        assert np.array_equal(tree._min, g[pre + "min"])


def test_heavy_tail_cap100000(golden):
    g = golden("tree_cap100000_heavy")
    rng = np.random.default_rng(int(g["seed"]))
    n = 100000
    fill_p = np.abs(rng.standard_normal(n)) * (1 + 9 * (rng.random(n) < 0.05)) + 1e-4
    tree = _replay_tape(g, fill_p=fill_p)
    assert np.array_equal(_sha(tree._sum), g["sum_sha"])
    assert np.array_equal(_sha(tree._min), g["min_sha"])
    assert np.array_equal(_sha(tree._val), g["val_sha"])
    # heavy tail => duplicates inside one stratified batch (last-writer-wins path)
    assert any(len(np.unique(r)) < len(r) for r in g["idx"])


def test_cfg1_per_sampler(golden):
    g = golden("per_cfg1_cap65536")
    cap, B = int(g["cap"]), int(g["batch"])
    s = po.OraclePERSampler(cap)
    # transform parity (per_sampler.py:16-17, float32 in) -- gate 1e-5 rel, expect ~ulp
    p = po.normalize_weights(g["fill_w"], 0.6)
    assert np.allclose(p, g["fill_p32"].astype(np.float64), rtol=1e-5, atol=0)
    # index parity is defined for identical leaf values: inject the reference's
    tail = 0
    for st in range(0, cap, 64):
        idx, tail = po.fifo_indices(cap, tail, 64)
        s.update(idx, g["fill_p32"][st:st + 64].astype(np.float64), 0.6, transform=False)
    for t in range(len(g["u"])):
        idx, w, _ = s.sample(B, g["u"][t], 0.4)
        assert np.array_equal(idx, g["idx"][t])
        assert np.allclose(w, g["w"][t], rtol=1e-12)
        assert s.sumtree.sum() == g["root"][t][0] and s.sumtree.min() == g["root"][t][1]
        pn = po.normalize_weights(g["neww"][t], 0.6)
        assert np.allclose(pn, g["newp32"][t].astype(np.float64), rtol=1e-5, atol=0)
        s.update(idx, g["newp32"][t].astype(np.float64), 0.6, transform=False)
    assert np.array_equal(_sha(s.sumtree._sum), g["sum_sha"])
    assert np.array_equal(_sha(s.sumtree._min), g["min_sha"])
    assert np.array_equal(_sha(s.sumtree._val), g["val_sha"])


def test_fifo_and_ring(golden):
    g = golden("fifo_ring")
    ring = po.OracleRing(10, [(np.uint8, (2, 3)), (np.int64, ()), (np.float32, ())])
    tail = 0
    for i, n in enumerate((4, 7, 3, 10, 1)):
        f, tail = po.fifo_indices(10, tail, n)
        assert np.array_equal(f, g[f"fifo{i}"])
        idx = ring.append_batch([g[f"in{i}_0"], g[f"in{i}_1"], g[f"in{i}_2"]])
        assert np.array_equal(idx, g[f"idx{i}"])
    out = ring.select(g["sel"])
    for c in range(3):
        assert np.array_equal(out[c], g[f"sel_{c}"])


def test_sampler_plugins_golden(golden):
    """Uniform / FIFO sampler restatements against the reference's own classes
    (sampler/uniform_sampler.py, sampler/fifo_sampler.py; make_golden.py section 9)."""
    g = golden("sampler_plugins")
    us, fs = po.OracleUniformSampler(300), po.OracleFIFOSampler(50)
    for k in range(5):
        us.update(g[f"msg{k}"], g[f"w{k}"])
        fs.update(g[f"msg{k}"], g[f"w{k}"])
        assert us.tail == int(g[f"u_tail{k}"])
        ui, uw = us.sample(g[f"u_draws{k}"])
        assert np.array_equal(ui, g[f"u_idx{k}"]) and np.array_equal(uw, g[f"u_w{k}"]) and uw.dtype == g[f"u_w{k}"].dtype
        if fs.ready_sample(8):
            fi, fw = fs.sample(8)
            assert np.array_equal(fi, g[f"f_idx{k}"]) and np.array_equal(fw, g[f"f_w{k}"])
        assert len(fs.buffer) == int(g[f"f_len{k}"])
    us.update(np.arange(5), np.ones(5))
    assert us.tail == int(g["u_tail_final"])
