"""DBoW2 vocabulary descent on the device against the CPU restatement of
TemplatedVocabulary::transform (oracle/port/camera.cc). Trees are synthetic (the reference's small_voc.yml.gz does not
travel to the GPU box): same shape as the shipped vocabulary (k = 9, L = 3) plus an uneven one with ties."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def make_tree(k, L, seed, uneven=False):
    """Nodes in a DBoW2-like file order: level by level below the root, then grandchildren appended parent by parent."""
    rng = np.random.default_rng(seed)
    node_id, parent_id, level = [], [], []
    frontier, nxt = [0], 1
    for lev in range(1, L + 1):
        new = []
        for p in frontier:
            kk = k if not uneven else int(rng.integers(1, k + 1))
            if uneven and lev > 1 and rng.uniform() < 0.25:
                continue            # p stays a leaf above the last level
            for _ in range(kk):
                node_id.append(nxt); parent_id.append(p); level.append(lev); new.append(nxt); nxt += 1
        frontier = new
    n = len(node_id)
    perm = np.arange(n)
    if uneven:                       # ids not ascending in file order (as nodeId 730 follows 9 in the shipped file)
        ids = rng.permutation(n) + 1
        remap = np.zeros(n + 1, np.int64); remap[1:] = ids
        node_id = [int(remap[i]) for i in node_id]; parent_id = [int(remap[p]) for p in parent_id]
    node_id, parent_id = np.array(node_id, np.int32), np.array(parent_id, np.int32)
    desc = rng.integers(0, 256, (n, 48), dtype=np.uint8)
    if uneven:                       # ties: some siblings share a descriptor -> the first child must win
        for i in range(1, n, 7):
            if parent_id[i] == parent_id[i - 1]:
                desc[i] = desc[i - 1]
    has_child = np.zeros(n + 1, bool); has_child[parent_id] = True
    leaf = ~has_child[node_id]
    word = np.full(n, -1, np.int32); word[leaf] = np.arange(int(leaf.sum()))
    weight = np.where(leaf, rng.uniform(0.0, 3.0, n), 0.0)
    weight[leaf & (rng.uniform(0, 1, n) < 0.05)] = 0.0      # stopped words have weight 0
    return node_id, parent_id, desc, weight, word


@pytest.mark.parametrize("k,L,uneven", [(9, 3, False), (10, 4, False), (6, 4, True), (2, 7, True)])
def test_vocabulary_transform_vs_port(k, L, uneven, mod, ctx, port, syn):
    tree = make_tree(k, L, 100 + k * L, uneven)
    voc = mod.Vocabulary(ctx, *tree, depth=L)
    rng = np.random.default_rng(7)
    feats = rng.integers(0, 256, (2500, 48), dtype=np.uint8)
    feats[:200] = tree[2][rng.integers(0, len(tree[2]), 200)]          # exact node descriptors: distance 0 paths
    for levelsup in (0, 1, 2, L, L + 3):
        w, wt, nid = voc.transform(feats, levelsup)
        pw, pwt, pnid = port.vocabulary_transform(*tree, L, feats, levelsup)
        assert np.array_equal(w, pw) and np.array_equal(nid, pnid), levelsup
        assert wt.tobytes() == pwt.tobytes(), levelsup
    if not uneven:
        assert len(np.unique(w)) > 20
    w0, _, _ = voc.transform(feats[:0])
    assert len(w0) == 0
    voc.close()


def test_vocabulary_create_rejects_malformed_trees(mod, ctx):
    nid, pid, desc, wt, word = make_tree(3, 2, 1)
    with pytest.raises(Exception):
        mod.Vocabulary(ctx, nid, pid, desc, wt, np.full(len(nid), -1, np.int32), depth=2)   # leaves without words
    bad = pid.copy(); bad[0] = nid[-1]; bad[-1] = nid[0]                                       # a cycle
    with pytest.raises(Exception):
        mod.Vocabulary(ctx, nid, bad, desc, wt, word, depth=2)
