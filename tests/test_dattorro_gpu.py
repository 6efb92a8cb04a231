"""GPU parity of the Dattorro plate bank through the C ABI.  The oracle for this row is
PINNED: bit-equal to libs/dattorro-verb/verb.cpp compiled unmodified (tests/test_oracle.py), so bit-exact here means
bit-exact against the reference itself; the committed golden vectors come from the compiled reference."""
import hashlib
import os

import numpy as np
import pytest
import torch

from ol_dsp_b200 import DattorroBank, dattorro
from tests.golden import make_golden as mg

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
NAN = float("nan")
SETTERS = [dattorro.PREDELAY, dattorro.PREFILTER, dattorro.INPUT_DIFFUSION1, dattorro.INPUT_DIFFUSION2, dattorro.DECAY,
           dattorro.DECAY_DIFFUSION, dattorro.DAMPING]


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def run_bank(x, params=None, chunks=None):
    """x: [frames][instances]; params: list (per instance) of 7-lists with NaN = keep default."""
    T, I = x.shape
    b = DattorroBank(I)
    if params is not None:
        for i, p in enumerate(params):
            for which, v in zip(SETTERS, p if p is not None else [NAN] * 7):
                if not np.isnan(v):
                    b.Set(which, float(v), first=i, count=1)
    dx = torch.from_numpy(np.ascontiguousarray(x)).cuda()
    dl, dr = torch.empty_like(dx), torch.empty_like(dx)
    a = 0
    for n in (chunks or [T]):
        b.ProcessDevice(dx[a:].data_ptr(), dl[a:].data_ptr(), dr[a:].data_ptr(), n)
        a += n
    torch.cuda.synchronize()
    b.close()
    return dl.cpu().numpy(), dr.cpu().numpy()


@pytest.mark.parametrize("name", list(mg.DATTORRO_CASES))
def test_bank_matches_reference_golden(orc, name):
    """Each golden case (outputs of the COMPILED REFERENCE) run as one instance of a small bank whose other instances
    carry different parameters and input: instances must not leak into each other."""
    g = np.load(os.path.join(G, "dattorro.npz"))
    params, seed, frames, scale = mg.DATTORRO_CASES[name]
    I = 5
    x = orc.noise(77, 0, I, 0, frames, 0.3)
    x[:, 2] = orc.noise(seed, 0, 1, 0, frames, scale)[:, 0]
    plist = [[0.5, 0.5, 0.5, 0.5, 0.5, 0.5, 0.5], None, params, [0.0, NAN, NAN, NAN, 0.9, NAN, NAN], [1.0, 0.1, NAN, NAN, NAN, NAN, NAN]]
    l, r = run_bank(x, plist, chunks=[frames // 3, frames - frames // 3])
    if f"{name}_l" in g:
        assert np.array_equal(bits(l[:, 2]), bits(g[f"{name}_l"])) and np.array_equal(bits(r[:, 2]), bits(g[f"{name}_r"]))
    else:
        assert [sha(np.ascontiguousarray(l[:, 2])), sha(np.ascontiguousarray(r[:, 2]))] == list(g[f"{name}_sha"])
    for i in (0, 3, 4):  # the neighbours against the oracle
        wl, wr = orc.dattorro(x[:, i], plist[i])
        assert np.array_equal(bits(l[:, i]), bits(wl)) and np.array_equal(bits(r[:, i]), bits(wr)), i


def test_wide_bank_ragged_chunks_and_reinit(orc):
    """300 instances (not a multiple of the block size), per-instance decay sweep as in C5, chunk sizes that straddle
    the 2 048-sample excursion steps, then Init() and a second run that must reproduce a fresh bank."""
    I, T = 300, 5000
    x = orc.noise(5, 100, I, 0, T, 0.25)
    decay = (0.5 + 0.4 * (np.arange(I) % 16) / 16).astype(np.float32)
    b = DattorroBank(I)
    b.SetDecay(decay)
    dx = torch.from_numpy(x).cuda()
    dl, dr = torch.empty_like(dx), torch.empty_like(dx)
    a = 0
    for n in (1, 2047, 1, 2000, 951):
        b.ProcessDevice(dx[a:].data_ptr(), dl[a:].data_ptr(), dr[a:].data_ptr(), n)
        a += n
    torch.cuda.synchronize()
    l, r = dl.cpu().numpy(), dr.cpu().numpy()
    for i in (0, 1, 31, 32, 150, 299):
        wl, wr = orc.dattorro(x[:, i], [NAN, NAN, NAN, NAN, float(decay[i]), NAN, NAN])
        assert np.array_equal(bits(l[:, i]), bits(wl)) and np.array_equal(bits(r[:, i]), bits(wr)), i
    b.Init()  # DattorroVerb_create again: defaults, cleared rings, t = 0
    b.ProcessDevice(dx.data_ptr(), dl.data_ptr(), dr.data_ptr(), 3000)
    torch.cuda.synchronize()
    wl, wr = orc.dattorro(x[:3000, 7], None)
    assert np.array_equal(bits(dl[:3000, 7].cpu().numpy()), bits(wl)) and np.array_equal(bits(dr[:3000, 7].cpu().numpy()), bits(wr))
    b.close()


def test_impulse_response_properties():
    """Size-independent properties at full width (8 192 instances, BASELINE configs[4]): silence in -> silence out;
    identical instances fed the same impulse give identical tails; the tail is non-zero and finite."""
    I, T = 8192, 6000
    b = DattorroBank(I)
    dx = torch.zeros((T, I), dtype=torch.float32, device="cuda")
    dl, dr = torch.empty_like(dx), torch.empty_like(dx)
    b.ProcessDevice(dx.data_ptr(), dl.data_ptr(), dr.data_ptr(), T)
    torch.cuda.synchronize()
    assert float(dl.abs().max()) == 0.0 and float(dr.abs().max()) == 0.0
    b.Init()
    dx[0, :] = 1.0
    b.ProcessDevice(dx.data_ptr(), dl.data_ptr(), dr.data_ptr(), T)
    torch.cuda.synchronize()
    assert torch.isfinite(dl).all() and torch.isfinite(dr).all()
    assert float(dl.abs().max()) > 0.0 and float(dr.abs().max()) > 0.0
    assert torch.equal(dl, dl[:, :1].expand(-1, I)) and torch.equal(dr, dr[:, :1].expand(-1, I))
    b.close()


@pytest.mark.parametrize("env", [{}, {"OLB_PLATE_KERNEL": "smem"}, {"OLB_PLATE_KERNEL": "split"}, {"OLB_PLATE_KERNEL": "rp"}, {"OLB_PLATE_KERNEL": "simple"}] + [{"OLB_PLATE_KERNEL": "pf", "OLB_PLATE_LANES": l} for l in ("32", "16", "8")],
                         ids=lambda e: "-".join(e.values()) or "default")
def test_every_kernel_variant_is_bit_exact(orc, monkeypatch, env):
    """The shared-memory-resident time-vectorised kernel (k_dattorro_smem, also what long blocks take by default), the
    kernel with the network cut over four warps, the one-warp ring-pointer kernel, the one-frame kernel and the
    index-arithmetic pipelined kernel at 32 / 16 / 8 active lanes per warp compute the same bits; chunks straddle the 2 048-frame excursion steps and the
    16-bit counter wrap is crossed (a first chunk of 63 000 silent frames)."""
    for k, v in env.items():
        monkeypatch.setenv(k, v)
    I, T = 77, 69000
    x = orc.noise(6, 3, I, 0, T, 0.25)
    x[:63000] = 0.0
    x[100:140] = orc.noise(7, 3, I, 0, 40, 0.25)
    plist = [[(i % 9) / 8.0, NAN, NAN, NAN, 0.5 + 0.4 * (i % 16) / 16, NAN, NAN] for i in range(I)]
    l, r = run_bank(x, plist, chunks=[2047, 3, 60950, 2500, T - 65500])
    for i in (0, 7, 8, 15, 16, 31, 32, 63, 64, 76):
        wl, wr = orc.dattorro(x[:, i], plist[i])
        assert np.array_equal(bits(l[:, i]), bits(wl)) and np.array_equal(bits(r[:, i]), bits(wr)), (env, i)


def test_smem_kernel_takes_long_blocks_and_hands_state_to_the_hbm_kernels(orc, monkeypatch):
    """Blocks of >= 8 192 frames run with the rings resident in shared memory (k_dattorro_smem), shorter ones with the rings
    in HBM (k_dattorro_split); both speak the same state format, so a stream that alternates between them is bit-exact."""
    monkeypatch.delenv("OLB_PLATE_KERNEL", raising=False)
    I, T = 37, 30000
    x = orc.noise(8, 3, I, 0, T, 0.25)
    plist = [[(i % 5) / 4.0, NAN, NAN, NAN, 0.5 + 0.4 * (i % 16) / 16, NAN, 0.3 + 0.02 * i] for i in range(I)]
    b = DattorroBank(I)
    for i, p in enumerate(plist):
        for which, v in zip(SETTERS, p):
            if not np.isnan(v):
                b.Set(which, float(v), first=i, count=1)
    dx = torch.from_numpy(x).cuda()
    dl, dr = torch.empty_like(dx), torch.empty_like(dx)
    a, used = 0, []
    for n in (9000, 1000, 8192, 11808):
        b.ProcessDevice(dx[a:].data_ptr(), dl[a:].data_ptr(), dr[a:].data_ptr(), n)
        used.append(b.LastKernel())
        a += n
    torch.cuda.synchronize()
    assert used == ["k_dattorro_smem", "k_dattorro_split", "k_dattorro_smem", "k_dattorro_smem"]
    l, r = dl.cpu().numpy(), dr.cpu().numpy()
    for i in (0, 1, 17, 36):
        wl, wr = orc.dattorro(x[:, i], plist[i])
        assert np.array_equal(bits(l[:, i]), bits(wl)) and np.array_equal(bits(r[:, i]), bits(wr)), i
    b.close()
