"""The *device* path of ShardedEmbedding end to end without a GPU.

tests/emu/sparse_emu_lib.cpp exposes psb_sparse_init / _stage / _pull / _apply with the C ABI of
libpsb_kernels.so, implemented by the device code of csrc/kernels/sparse.cu run thread by thread
on the host.  Loaded in place of the real library (and with torch.cuda's stream / device context
calls stubbed), it lets the Python side of the GPU path run on host tensors: the ctypes PsbSparse
struct, the peer-pointer tables, argument order, staging, barriers-free emulated ranks, the
resting state of the scratch.  It is tests/test_zy_sparse_gpu.py minus the GPU; what remains for
the GPU is the hardware (alignment faults, vector atomics, NVLink addressing)."""
import contextlib
import ctypes as C
import shutil
import subprocess
import types
from pathlib import Path

import pytest
import torch

from parameter_server_distributed_b200 import OptimConfig, _lib
from parameter_server_distributed_b200.parallel import sparse
from parameter_server_distributed_b200.parallel.sparse import ShardedEmbedding

ROOT = Path(__file__).resolve().parents[1]
pytestmark = pytest.mark.skipif(shutil.which("g++") is None, reason="needs g++")


def _cut(text, start, end):
    a = text.index(start)
    return text[a:text.index(end, a)]


@pytest.fixture(scope="module")
def emulated_lib(tmp_path_factory):
    d = tmp_path_factory.mktemp("emu_lib")
    (d / "sparse_kernels.inc").write_text(_cut((ROOT / "csrc/kernels/sparse.cu").read_text(),
                                               "constexpr int kThreads", "// ---- host side"))
    (d / "optim_math.inc").write_text(_cut((ROOT / "csrc/kernels/common.cuh").read_text(),
                                           "struct OptimCoef {", "}  // namespace psb"))
    so = d / "libsparse_emu.so"
    cc = subprocess.run(["g++", "-std=c++17", "-O1", "-shared", "-fPIC", "-I", str(d), "-I",
                         str(ROOT / "tests/emu"), "-I", str(ROOT / "include"),
                         str(ROOT / "tests/emu/sparse_emu_lib.cpp"), "-o", str(so)],
                        capture_output=True, text=True)
    assert cc.returncode == 0, cc.stderr[-4000:]
    lib = C.CDLL(str(so))
    vp, i64 = C.c_void_p, C.c_int64
    lib.psb_last_error.restype = C.c_char_p
    for name, args in {"psb_sparse_init": [vp, i64, vp],
                       "psb_sparse_stage": [C.POINTER(_lib.PsbSparse), vp, vp, C.c_int, i64, vp],
                       "psb_sparse_pull": [C.POINTER(_lib.PsbSparse), vp, i64, vp, vp],
                       "psb_sparse_apply": [C.POINTER(_lib.PsbSparse), vp],
                       "psb_counter_add": [vp, C.c_uint32, vp]}.items():
        getattr(lib, name).restype = C.c_int
        getattr(lib, name).argtypes = args
    return lib


@pytest.fixture()
def device_path(emulated_lib, monkeypatch):
    """Context in which ShardedEmbedding(...) objects take the device path on host tensors."""
    monkeypatch.setattr(_lib, "load", lambda build_if_missing=True: emulated_lib)
    monkeypatch.setattr(torch.cuda, "device", lambda dev: contextlib.nullcontext())
    stream = types.SimpleNamespace(cuda_stream=0, synchronize=lambda: None)
    monkeypatch.setattr(torch.cuda, "current_stream", lambda dev=None: stream)

    @contextlib.contextmanager
    def on():
        sparse._DEVICE_PATH_ON_HOST = True
        try:
            yield
        finally:
            sparse._DEVICE_PATH_ON_HOST = False
    return on


class _Buf:
    def __init__(self, group, rank, k):
        self.group, self.rank, self.k = group, rank, k

    @property
    def local(self):
        return self.group.bufs[self.rank][self.k]

    def view(self, dtype):
        return self.local.view(dtype)

    def peer_ptrs(self):
        return [self.group.bufs[w][self.k].data_ptr() for w in range(self.group.world)]

    def free(self):
        pass


class _Group:
    """Symmetric allocations of several emulated ranks in host memory (Fabric.local_group's role)."""

    def __init__(self, world):
        self.world = world
        self.bufs = [[] for _ in range(world)]
        self.count = [0] * world

    def fabric(self, rank):
        g = self

        def alloc(nbytes, align=256):
            k = g.count[rank]
            g.count[rank] += 1
            if k == len(g.bufs[0]):  # first rank to ask allocates the range on every rank
                for w in range(g.world):
                    g.bufs[w].append(torch.zeros(nbytes + 64, dtype=torch.uint8)[:nbytes])
            assert g.bufs[rank][k].numel() == nbytes
            return _Buf(g, rank, k)
        return types.SimpleNamespace(rank=rank, world=self.world, device=torch.device("cpu"),
                                     alloc=alloc, _owns=False,
                                     c_fabric=lambda: _lib.PsbFabric(rank, g.world))


def _batches(world, n, vocab, dim, steps, seed):
    gen = torch.Generator().manual_seed(seed)
    out = []
    for _ in range(steps):
        idx = torch.randint(0, vocab, (world, n), generator=gen)
        idx[:, :n // 8] = idx[0, :n // 8].clone()
        idx[:, n // 8:n // 4] = idx[:, n // 4:n // 4 + n // 8].clone()
        idx[:, -1] = -1
        out.append((idx, torch.randn(world, n, dim, generator=gen)))
    return out


@pytest.mark.parametrize("kind,dim,gdt", [("sgd", 16, torch.float32), ("adam", 128, torch.bfloat16),
                                          ("momentum", 264, torch.bfloat16)])
def test_one_rank_device_path_matches_the_host_path(device_path, kind, dim, gdt):
    vocab, n = 1203, 128
    cfg = OptimConfig(kind=kind, lr=0.05)
    with device_path():
        dev = ShardedEmbedding(vocab, dim, cfg, capacity=n, grad_dtype=gdt, seed=4)
    host = ShardedEmbedding(vocab, dim, cfg, capacity=n, grad_dtype=gdt, seed=4)
    assert dev.cuda and not host.cuda and torch.equal(dev.master, host.master)
    assert torch.equal(dev.table, host.table) and bool((dev.slot == -1).all())
    for idx, g in _batches(1, n, vocab, dim, 3, seed=11):
        send = g[0].to(torch.bfloat16) if kind == "adam" else g[0]   # both input dtypes of stage
        dev.push(idx[0], send)
        host.push(idx[0], send)
    assert (dev.master - host.master).abs().max() < 1e-5
    assert int(dev.step_word.item()) == 3 == dev.steps
    q = torch.cat([torch.randint(0, vocab, (100,)), torch.tensor([-1, vocab, 0, vocab - 1])])
    got, want = dev.pull(q), host.pull(q)
    assert (got.float() - want.float()).abs().max() < 1e-3   # bf16 of masters within 1e-5
    assert torch.equal(got[102], dev.master[0].to(torch.bfloat16))
    assert got[100].abs().sum() == 0 and got[101].abs().sum() == 0
    assert int(dev.n_touched.item()) == 0 and bool((dev.slot == -1).all())
    assert float(dev.acc.abs().max()) == 0.0
    # the autograd front end and the checkpoint on the same path
    loss = dev.lookup(torch.tensor([[3, 4], [3, 9]])).float().sum()
    loss.backward()
    before = dev.master.clone()
    dev.apply_gradients()
    assert (dev.master[[3, 4, 9]] - before[[3, 4, 9]]).abs().min() > 0
    sd = dev.state_dict()
    with device_path():
        again = ShardedEmbedding(vocab, dim, cfg, capacity=n, grad_dtype=gdt, seed=99)
    again.load_state_dict(sd)
    assert torch.equal(again.master, dev.master) and torch.equal(again.table, dev.table)
    assert int(again.step_word.item()) == dev.steps


def test_four_emulated_ranks_on_the_device_path(device_path):
    world, vocab, dim, n = 4, 1001, 64, 96
    cfg = OptimConfig(kind="adamw", lr=0.02, weight_decay=0.05)
    group = _Group(world)
    with device_path():
        embs = [ShardedEmbedding(vocab, dim, cfg, capacity=n, fabric=group.fabric(r), seed=7)
                for r in range(world)]
    host = ShardedEmbedding(vocab, dim, cfg, capacity=n * world, seed=7, grad_dtype=torch.float32)
    assert [e.rank for e in embs] == [0, 1, 2, 3] and all(e._emulated for e in embs)
    for idx, g in _batches(world, n, vocab, dim, 3, seed=12):
        for r, e in enumerate(embs):
            e.stage(idx[r], g[r])
        for e in embs:
            e.apply()
        gb = g.to(torch.bfloat16).float()
        host.push(idx.reshape(-1), (gb / world).reshape(-1, dim))
    for e in embs:
        own = (e.master[:e.hi - e.lo] - host.master[e.lo:e.hi]).abs()
        assert (own > 2e-4).float().mean() < 1e-4 and own.max() < 0.2
    want = host.table.float()
    for e in (embs[0], embs[3]):
        diff = (e.full_table().float() - want).abs()
        assert (diff > 1e-3).float().mean() < 1e-4
    with pytest.raises(RuntimeError, match="emulated ranks"):
        embs[0].push(torch.arange(4), torch.zeros(4, dim))
