"""Two-rank NCCL refit: each rank holds half the rows, the only exchange is the all-reduce of the
[loss | grad] vector, and both ranks must land on the single-rank model.  Needs 2 GPUs (skipped on
a 1-GPU box; run with `gpurun --gpus 2`)."""
import os
import socket

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _problem(kind):
    rng = np.random.default_rng(4)
    if kind == "binary_d32":                       # fused lr_loss_grad
        n, d = 40_001, 32
        X = rng.standard_normal((n, d), dtype=np.float32)
        y = (X[:, 0] * 1.5 - X[:, 3] + 0.3 * rng.standard_normal(n) > 0.2).astype(np.int64)
        return X, y, np.array([0, 1])
    n, d, c = 8_001, 64, 20                        # general path: tensor-core forward pass, 10.4 KB all-reduce
    X = rng.standard_normal((n, d), dtype=np.float32)
    W = rng.standard_normal((c, d)) * 0.5
    y = np.argmax(X @ W.T + rng.gumbel(size=(n, c)), axis=1).astype(np.int64)
    return X, y, np.arange(c)


def _worker(rank, world, port, q, kind="binary_d32"):
    import torch
    import torch.distributed as dist
    from mlapi_b200 import refit, sharding
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world,
                            device_id=torch.device("cuda", rank))
    X, y, classes = _problem(kind)
    lo, hi = sharding.my_rows(len(X), rank, world)
    res = refit.fit(X[lo:hi], y[lo:hi], classes=classes, max_iter=200)
    q.put((rank, res.coef_, res.intercept_, res.n_iter_[0], res.world_))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("kind", ["binary_d32", "multinomial_20"])
def test_two_rank_refit_matches_single_rank(kind):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import __graft_entry__
    __graft_entry__.build()
    import torch.multiprocessing as mp
    from mlapi_b200 import refit
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q, kind)) for r in range(2)]
    [p.start() for p in procs]
    results = sorted((q.get(timeout=300) for _ in procs), key=lambda t: t[0])
    [p.join(timeout=60) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    X, y, classes = _problem(kind)
    single = refit.fit(X, y, classes=classes, max_iter=200)
    assert results[0][4] == 2
    np.testing.assert_array_equal(results[0][1], results[1][1])          # identical steps on both ranks
    tol = 2e-3 if kind == "binary_d32" else 2e-2                          # both stop at lbfgs's gtol = 1e-4
    np.testing.assert_allclose(results[0][1], single.coef_, atol=tol)
    np.testing.assert_allclose(results[0][2], single.intercept_, atol=tol)


def test_estimator_shards_host_batches_across_two_gpus():
    """`devices=[0, 1]`: contiguous row blocks, one replica and one host thread per GPU, no collective;
    the reassembled result is bit-identical to the single-GPU one."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import __graft_entry__
    __graft_entry__.build()
    import sys
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    import synth
    from mlapi_b200 import B200LogisticRegression
    X, W, b = synth.make_config("cfg1_d32_binary", 300_001)
    one = B200LogisticRegression(W, b, [0, 1], devices=[0])
    two = B200LogisticRegression(W, b, [0, 1], devices=[0, 1])
    a = one.score_rows(X, proba=True, label=True, decision=True, maxprob=True)
    c = two.score_rows(X, proba=True, label=True, decision=True, maxprob=True)
    for k in a:
        assert np.array_equal(a[k], c[k]), k
    one.close()
    two.close()
