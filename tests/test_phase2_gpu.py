"""GPU parity for kernel (4): rcq_phase2_host against the reference-generated fixtures (bit-exact
valid set, integer sums, fp64 matrix) and against the numpy oracle on larger random inputs."""
import numpy as np
import pytest

from test_phase2_oracle import CASES, load_case

pytestmark = pytest.mark.gpu


def run_gpu(ctx, n_samples, tables, sample_count=None):
    from rna_clique_b200 import _native
    stride = 1 + max([0] + [int(max(t["sgene"].max(initial=0), t["qgene"].max(initial=0))) for t in tables])
    row_off = np.concatenate([[0], np.cumsum([len(t["sgene"]) for t in tables])]).astype(np.int64)
    cat = lambda c: np.concatenate([t[c] for t in tables]) if tables else np.zeros(0, dtype=np.int32)
    res = _native.phase2_host(ctx, n_samples, stride, row_off, [t["s"] for t in tables], [t["q"] for t in tables],
                              cat("sgene"), cat("qgene"), cat("nident"), cat("length"), cat("gaps"), sample_count)
    res["stride"] = stride
    return res


@pytest.mark.parametrize("name", CASES)
def test_phase2_gpu_against_reference(ctx, name):
    from oracle import phase2_ref
    rec, names, tables = load_case(name)
    exp = rec["expect"]
    res = run_gpu(ctx, len(names), tables, rec.get("sample_count"))
    assert res["counts"]["sample_count"] == exp["sample_count"]
    assert res["counts"]["nodes"] == exp["n_nodes"] and res["counts"]["edges"] == exp["n_edges"]
    v = np.flatnonzero(res["valid"])
    got_valid = sorted([names[int(i // res["stride"])], int(i % res["stride"])] for i in v)
    assert got_valid == exp["valid"]
    assert res["no_ideal"] == bool(exp["error"])
    if exp["error"]:
        return
    sums = list(zip(res["nident_sum"].tolist(), res["denom_sum"].tolist()))
    labels, m = phase2_ref.distance_matrix(names, tables, sums)
    assert labels == exp["samples"]
    assert np.array_equal(m, np.asarray(exp["matrix"]))


def test_phase2_gpu_random_large(ctx):
    from oracle import phase2_ref
    rng = np.random.default_rng(3)
    n_samples, n_genes = 12, 4000
    tables = []
    for i in range(n_samples):
        for j in range(i + 1, n_samples):
            g = np.flatnonzero(rng.random(n_genes) > 0.01).astype(np.int32)
            q = g.copy()
            wrong = rng.random(len(g)) < 0.003
            q[wrong] = rng.integers(0, n_genes, size=int(wrong.sum()))
            dup = rng.random(len(g)) < 0.4
            sg, qg = np.concatenate([g, g[dup]]), np.concatenate([q, q[dup]])
            length = rng.integers(300, 3000, size=len(sg)).astype(np.int32)
            gaps = rng.integers(0, 4, size=len(sg)).astype(np.int32)
            nident = (length - gaps - rng.integers(0, 50, size=len(sg))).astype(np.int32)
            p = rng.permutation(len(sg))
            tables.append(dict(s=i, q=j, sgene=sg[p], qgene=qg[p], nident=nident[p], length=length[p], gaps=gaps[p]))
    exp = phase2_ref.phase2(n_samples, tables)
    res = run_gpu(ctx, n_samples, tables)
    assert res["counts"]["ideal"] == exp["n_ideal"] and exp["n_ideal"] > 100
    # the four numbers of `plot_component_sizes --statistics` (plot_component_sizes.py:214-230)
    assert (res["counts"]["sample_count"], res["counts"]["components"], res["counts"]["ge_samples"]) == \
        (exp["sample_count"], exp["n_components"], exp["n_ge_samples"])
    assert exp["n_ge_samples"] > exp["n_ideal"]
    assert res["counts"]["nodes"] == exp["n_nodes"] and res["counts"]["edges"] == exp["n_edges"]
    got = {(int(i // res["stride"]), int(i % res["stride"])) for i in np.flatnonzero(res["valid"])}
    assert got == exp["valid"]
    assert list(zip(res["nident_sum"].tolist(), res["denom_sum"].tolist())) == exp["sums"]


def test_phase2_empty_and_single(ctx):
    res = run_gpu(ctx, 2, [])
    assert res["counts"]["nodes"] == 0 and not res["no_ideal"]
    t = dict(s=0, q=1, sgene=np.array([5], dtype=np.int32), qgene=np.array([7], dtype=np.int32),
             nident=np.array([90], dtype=np.int32), length=np.array([100], dtype=np.int32), gaps=np.array([2], dtype=np.int32))
    res = run_gpu(ctx, 2, [t])
    assert res["nident_sum"].tolist() == [90] and res["denom_sum"].tolist() == [98] and res["counts"]["ideal"] == 1
