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


def _worker(rank, world, port, dims, acts, rows, micro, steps, out_dir):
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
    ps = PipelineStage(stage, rank, world)
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


@pytest.mark.parametrize("world,micro", [(2, 4), (3, 5)])
def test_gpipe_schedule_equals_full_batch_oracle(tmp_path, world, micro):
    import oracle
    from ddnn_b200.config import make_params, synthetic_batch
    dims, acts = [12, 10, 9, 8, 5], ["relu", "sigmoid", "relu", "softmax"]
    rows, steps = 22, 3  # 22 rows / 4 or 5 micro-batches: ragged micro-batches
    mp.spawn(_worker, args=(world, _free_port(), dims, acts, rows, micro, steps, str(tmp_path)), nprocs=world,
             join=True)
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
