"""The pipeline schedule / protocol on CPU: world_size 2 and 3 over gloo, with a numpy stage
backed by the float64 oracle standing in for the GPU ``Layer`` (same duck-typed interface).
Checks that GPipe micro-batching with the deferred full-batch dW step reproduces the
single-process full-batch oracle step exactly."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class OracleStage:
    """CPU stand-in for ddnn_b200.Layer's training interface (tests only)."""

    def __init__(self, params, in_dim, input_activation, is_output):
        import oracle
        self.o = oracle
        self.params = params
        self.expected_input_dim = in_dim
        self.input_activation = input_activation
        self.is_output = is_output

    def begin_batch(self, rows):
        p = self.params
        self.a = [torch.zeros((rows, p[0]["weights"].shape[0]), dtype=torch.float64)]
        self.z, self.dz = [], []
        for q in p:
            n = q["weights"].shape[1]
            self.a.append(torch.zeros((rows, n), dtype=torch.float64))
            self.z.append(torch.zeros((rows, n), dtype=torch.float64))
            self.dz.append(torch.zeros((rows, n), dtype=torch.float64))
        self.dx = torch.zeros_like(self.a[0])
        self.loss = torch.zeros(1, dtype=torch.float64)

    def input_buffer(self, r0, n):
        return self.a[0][r0:r0 + n]

    def grad_buffer(self, r0, n):
        return self.dz[-1][r0:r0 + n]

    def forward_train(self, x, r0=0, rows=None):
        if x is not None:
            rows = x.shape[0]
            self.a[0][r0:r0 + rows] = x
        acts, zs = self.o.forward_all(self.params, self.a[0][r0:r0 + rows].numpy())
        for i in range(len(self.params)):
            self.a[i + 1][r0:r0 + rows] = torch.from_numpy(acts[i + 1])
            self.z[i][r0:r0 + rows] = torch.from_numpy(zs[i])
        return self.a[-1][r0:r0 + rows]

    def loss_and_grad(self, labels, r0=0, denom=None, accumulate=False):
        n = labels.numel()
        l, _, dz = self.o.softmax_ce(self.z[-1][r0:r0 + n].numpy(), labels.numpy(), denom)
        self.dz[-1][r0:r0 + n] = torch.from_numpy(dz)
        self.loss = (self.loss if accumulate else 0) + torch.tensor([l], dtype=torch.float64)
        return self.loss

    def backward(self, r0=0, rows=None, need_input_grad=True):
        ab = self.o.ddnn_oracle._act_backward
        sl = slice(r0, r0 + rows)
        for i in range(len(self.params) - 1, 0, -1):
            da = self.dz[i][sl].numpy() @ self.params[i]["weights"].T
            self.dz[i - 1][sl] = torch.from_numpy(ab(da, self.a[i][sl].numpy(), self.z[i - 1][sl].numpy(),
                                                     self.params[i - 1]["activation"]))
        if not need_input_grad:
            return None
        da = self.dz[0][sl].numpy() @ self.params[0]["weights"].T
        a0 = self.a[0][sl].numpy()
        # derivative of the upstream stage's activation from its OUTPUT (what the GPU Layer does)
        g = {"relu": (a0 > 0).astype(np.float64), "sigmoid": a0 * (1 - a0)}.get(self.input_activation, 1.0)
        self.dx[sl] = torch.from_numpy(da * g)
        return self.dx[sl]

    def step(self, lr, rows=None):
        for i, p in enumerate(self.params):
            p["weights"] = p["weights"] - lr * (self.a[i][:rows].numpy().T @ self.dz[i][:rows].numpy())
            p["biases"] = p["biases"] - lr * self.dz[i][:rows].numpy().sum(axis=0)


    def accumulate_grads(self, r0, r1, first):
        if first:
            self.gw = [np.zeros_like(p["weights"]) for p in self.params]
            self.gb = [np.zeros_like(p["biases"]) for p in self.params]
        for i in range(len(self.params)):
            self.gw[i] += self.a[i][r0:r1].numpy().T @ self.dz[i][r0:r1].numpy()
            self.gb[i] += self.dz[i][r0:r1].numpy().sum(axis=0)

    def apply_update(self, lr):
        for i, p in enumerate(self.params):
            p["weights"] = p["weights"] - lr * self.gw[i]
            p["biases"] = p["biases"] - lr * self.gb[i]


def _worker(rank, world, port, dims, acts, rows, micro, steps, out_dir, schedule="gpipe", dw_chunk=None):
    sys.path.insert(0, REPO)
    sys.path.insert(0, os.path.join(REPO, "tests"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from ddnn_b200.config import make_params, synthetic_batch
    from ddnn_b200.pipeline import PipelineStage, stage_assignment
    params = make_params(dims, acts, seed=1)
    lo, hi = stage_assignment(len(acts), world)[rank]
    stage = OracleStage(params[lo:hi], dims[lo], acts[lo - 1] if lo > 0 else "linear", hi == len(acts))
    ranges = stage_assignment(len(acts), world)
    ps = PipelineStage(stage, rank, world, schedule=schedule, dw_chunk=dw_chunk,
                       layers_per_stage=[b - a for a, b in ranges])
    losses = []
    for s in range(steps):
        x, y = synthetic_batch(rows, dims[0], dims[-1], seed=s)
        xt = torch.from_numpy(x) if rank == 0 else None
        yt = torch.from_numpy(y) if rank == world - 1 else None
        l = ps.train_step(xt, yt, rows, micro, 0.1)
        if l is not None:
            losses.append(float(l.item()))
    np.savez(os.path.join(out_dir, f"rank{rank}.npz"), losses=np.array(losses),
             **{f"w{lo + i}": p["weights"] for i, p in enumerate(stage.params)},
             **{f"b{lo + i}": p["biases"] for i, p in enumerate(stage.params)})
    dist.barrier()
    dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("world,micro,schedule,dw_chunk", [(2, 4, "gpipe", None), (3, 5, "gpipe", None),
                                                           (3, 5, "gpipe", 2), (3, 6, "zb", 2), (2, 5, "zb", 1)])
def test_pipeline_schedules_equal_full_batch_oracle(tmp_path, world, micro, schedule, dw_chunk):
    """GPipe with the deferred full-batch dW, GPipe with chunked gradient accumulation and the zero-bubble order
    (forwards / backwards interleaved, weight-gradient chunks in the waits) all reproduce the single-process
    full-batch oracle step (float64: the chunked partial sums differ from the one-piece sum by ~1e-16)."""
    import oracle
    from ddnn_b200.config import make_params, synthetic_batch
    dims, acts = [12, 10, 9, 8, 5], ["relu", "sigmoid", "relu", "softmax"]
    rows, steps = 22, 3  # 22 rows / 4, 5 or 6 micro-batches: ragged micro-batches
    mp.spawn(_worker, args=(world, _free_port(), dims, acts, rows, micro, steps, str(tmp_path), schedule, dw_chunk),
             nprocs=world, join=True)
    ref = make_params(dims, acts, seed=1)
    ref_losses = []
    for s in range(steps):
        x, y = synthetic_batch(rows, dims[0], dims[-1], seed=s)
        ref_losses.append(oracle.train_step(ref, x, y, 0.1))
    got = {}
    for r in range(world):
        got.update(dict(np.load(os.path.join(str(tmp_path), f"rank{r}.npz"))))
    np.testing.assert_allclose(got["losses"], ref_losses, rtol=1e-12)
    for i, p in enumerate(ref):
        np.testing.assert_allclose(got[f"w{i}"], p["weights"], rtol=1e-10, atol=1e-14)
        np.testing.assert_allclose(got[f"b{i}"], p["biases"], rtol=1e-10, atol=1e-14)


@pytest.mark.parametrize("lps,micro,chunk", [([1] * 8, 32, 8), ([2, 2, 2, 2], 32, 4), ([4, 4], 16, 2), ([2, 1, 1], 6, 2),
                                             ([3], 5, 2)])
def test_zero_bubble_order_is_complete_and_consistent(lps, micro, chunk):
    """Every stage's list holds each F / B / W op once and the update last; F(m) never precedes what it needs from
    the stage before, B(m) what it needs from the stage after, W(c) the B ops of its chunk; replaying the lists
    with unit costs and in-order stages terminates (no deadlock) no later than the GPipe order."""
    from ddnn_b200.pipeline import pipeline_schedule
    S = len(lps)
    nchunks = -(-micro // chunk)

    def replay(order):
        fd = [[None] * micro for _ in range(S)]
        bd = [[None] * micro for _ in range(S)]
        ptr, free = [0] * S, [0.0] * S
        progress = True
        while progress:
            progress = False
            for i in range(S):
                while ptr[i] < len(order[i]):
                    k, j = order[i][ptr[i]]
                    if k == "F":
                        src = 0.0 if i == 0 else fd[i - 1][j]
                        if src is None:
                            break
                        free[i] = max(free[i], src) + lps[i]
                        fd[i][j] = free[i]
                    elif k == "B":
                        src = fd[i][j] if i == S - 1 else bd[i + 1][j]
                        if src is None or fd[i][j] is None:
                            break
                        free[i] = max(free[i], src) + (lps[i] - 1 if i == 0 else lps[i])
                        bd[i][j] = free[i]
                    elif k == "W":
                        ms = range(micro) if j < 0 else range(j * chunk, min(micro, (j + 1) * chunk))
                        if any(bd[i][m] is None for m in ms):
                            break
                        free[i] = max([free[i]] + [bd[i][m] for m in ms]) + lps[i] * len(ms)
                    ptr[i] += 1
                    progress = True
        assert all(ptr[i] == len(order[i]) for i in range(S)), "the op lists deadlock"
        return max(free)

    zb = pipeline_schedule("zb", lps, micro, chunk)
    for ops in zb:
        assert sorted(o for o in ops if o[0] == "F") == [("F", m) for m in range(micro)]
        assert sorted(o for o in ops if o[0] == "B") == [("B", m) for m in range(micro)]
        assert sorted(o for o in ops if o[0] == "W") == [("W", c) for c in range(nchunks)]
        assert ops[-1] == ("U", 0) and len(ops) == 2 * micro + nchunks + 1
        assert [m for k, m in ops if k == "F"] == list(range(micro))  # forwards stay in order (loss accumulation too)
        assert [m for k, m in ops if k == "B"] == list(range(micro))
    t_zb = replay(zb)
    t_gpipe = replay(pipeline_schedule("gpipe", lps, micro, chunk))
    assert t_zb <= t_gpipe
    if S == 8:
        assert t_zb == 3 * micro + (S - 1)  # the last stage's bound: (S - 1) fill slots + its own 3M of work
    with pytest.raises(ValueError):
        pipeline_schedule("zb", lps, micro, None)
