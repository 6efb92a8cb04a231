"""GPU parity of the biquad / one-pole cascade bank against the oracle.
Everything goes through the C ABI (ctypes); bit-exact is the bar in BOTH modes, because the GPU FMA
kernels perform the same fused operations in the same order as the oracle's FMA mode."""
import numpy as np
import pytest
import torch

from oracle import loader
from ol_dsp_b200 import (KIND_LPF, KIND_PORT, LAYOUT_FRAMES, LAYOUT_PLANAR, MODE_EXACT, MODE_FAST, MODE_FMA, BiquadBank, OlbError,
                         _lib, c2_coefs, noise_device)

pytestmark = pytest.mark.gpu

V_AUTO, V_DIRECT, V_TMA, V_SYS, V_SKEW = 0, 1, 2, 3, 4
KNAME = {V_DIRECT: "k_bq_direct", V_TMA: "k_bq_tma", V_SYS: "k_bq_sys", V_SKEW: "k_bq_skew"}


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def run_bank(x, k, mode, variant=V_AUTO, kinds=None, layout=LAYOUT_FRAMES, chunks=None):
    C = k.shape[0]
    b = BiquadBank(C, k.shape[1])
    if kinds is not None:
        b.SetKinds(kinds)
    b.SetMode(mode)
    b.SetVariant(variant)
    b.SetCoeffs(k)
    if chunks is None:
        y = b.ProcessBlock(x, layout)
    else:
        parts, a = [], 0
        for n in chunks:
            parts.append(b.ProcessBlock(x[a:a + n] if layout == LAYOUT_FRAMES else np.ascontiguousarray(x[:, a:a + n]), layout))
            a += n
        y = np.concatenate(parts, axis=0 if layout == LAYOUT_FRAMES else 1)
    name, st = b.LastKernel(), b.GetState()
    b.close()
    return y, name, st


def test_noise_matches_oracle(orc):
    for layout, shape in ((LAYOUT_FRAMES, (1003, 37)), (LAYOUT_PLANAR, (37, 1003))):
        t = torch.empty(shape, dtype=torch.float32, device="cuda")
        noise_device(t.data_ptr(), 99, 5, 37, 1234567, 1003, 0.5, layout)
        torch.cuda.synchronize()
        ref = orc.noise(99, 5, 37, 1234567, 1003, 0.5, planar=layout == LAYOUT_PLANAR)
        assert np.array_equal(bits(t.cpu().numpy()), bits(ref))


@pytest.mark.parametrize("mode", [MODE_EXACT, MODE_FMA])
@pytest.mark.parametrize("variant,C", [(V_DIRECT, 256), (V_DIRECT, 77), (V_TMA, 256), (V_TMA, 68), (V_TMA, 4),
                                       (V_SYS, 256), (V_SYS, 68), (V_SYS, 4), (V_SYS, 9600)])
def test_c2_cascade_bit_exact(orc, mode, variant, C):
    T = 2051  # ragged: not a multiple of any tile height
    x = orc.noise(1, 0, C, 0, T)
    k = c2_coefs(C)
    y, name, st = run_bank(x, k, mode, variant)
    assert name == KNAME[variant]
    ost = np.zeros((C, 8, 4), dtype=np.float32)
    want = orc.biquad(x, k, mode=mode, state=ost)
    assert np.array_equal(bits(y), bits(want))
    assert np.array_equal(bits(st), bits(ost))


@pytest.mark.parametrize("S", [1, 2, 3, 5, 7, 8, 12, 16])
def test_section_counts(orc, S):
    C, T = 128, 700
    x = orc.noise(2, 0, C, 0, T)
    k = np.ascontiguousarray(np.tile(c2_coefs(C), (1, 2, 1))[:, :S])
    for mode in (MODE_EXACT, MODE_FMA):
        y, name, _ = run_bank(x, k, mode)
        assert np.array_equal(bits(y), bits(orc.biquad(x, k, mode=mode))), (S, mode, name)


@pytest.mark.parametrize("mode", [MODE_EXACT, MODE_FMA])
def test_one_pole_kinds_reproduce_port_and_verb_lpf(orc, mode):
    """PORT = daisysp::Port (Portamento.h:29-32), LPF = LowPassFilter_process (verb.cpp:131-134)."""
    C, T = 96, 1500
    x = orc.noise(3, 0, C, 0, T)
    kinds = np.array([KIND_PORT, KIND_LPF, 0, 0, KIND_LPF, KIND_PORT, 0, 0], dtype=np.uint8)
    k = c2_coefs(C)
    k[:, [0, 5], 0] = np.linspace(0.5, 0.999, C, dtype=np.float32)[:, None]
    k[:, [1, 4], 0] = np.linspace(0.05, 0.95, C, dtype=np.float32)[:, None]
    ost = np.zeros((C, 8, 4), dtype=np.float32)
    want = orc.biquad(x, k, kinds=kinds, mode=mode, state=ost)
    y, name, st = run_bank(x, k, mode, kinds=kinds)
    assert np.array_equal(bits(y), bits(want)), name
    assert np.array_equal(bits(st), bits(ost))


@pytest.mark.parametrize("variant", [V_DIRECT, V_TMA, V_SYS, V_SKEW])
def test_state_carries_across_calls_and_per_frame_process(orc, variant):
    C, T = 64, 300
    x = orc.noise(4, 0, C, 0, T)
    k = c2_coefs(C)
    want = orc.biquad(x, k, mode=MODE_FMA)
    y, _, _ = run_bank(x, k, MODE_FMA, variant, chunks=[1, 1, 1, 31, 32, 33, 201])
    assert np.array_equal(bits(y), bits(want))


def test_planar_layout(orc):
    C, T = 70, 1111
    x = orc.noise(6, 0, C, 0, T, planar=True)
    k = c2_coefs(C)
    for mode in (MODE_EXACT, MODE_FMA):
        y, name, _ = run_bank(x, k, mode, layout=LAYOUT_PLANAR, chunks=[500, 611])
        assert name == "k_bq_planar"
        assert np.array_equal(bits(y), bits(orc.biquad(x, k, mode=mode, planar=True)))


def test_init_resets_state_and_set_state_round_trip(orc):
    C, T = 64, 256
    x = orc.noise(7, 0, C, 0, T)
    k = c2_coefs(C)
    b = BiquadBank(C, 8)
    b.SetMode(MODE_EXACT)
    b.SetCoeffs(k)
    y1 = b.ProcessBlock(x)
    st = b.GetState()
    y2 = b.ProcessBlock(x)
    b.Init(48000.0)  # re-Init resets state (test/fx_test.cpp:32,42)
    assert not b.GetState().any()
    assert np.array_equal(bits(b.ProcessBlock(x)), bits(y1))
    b.SetState(st)
    assert np.array_equal(bits(b.ProcessBlock(x)), bits(y2))
    b.close()


def test_setter_takes_effect_at_next_block_for_a_channel_range(orc):
    C, T = 128, 200
    x = orc.noise(8, 0, C, 0, 2 * T)
    k = c2_coefs(C)
    b = BiquadBank(C, 8)
    b.SetMode(MODE_FMA)
    b.SetCoeffs(k)
    ost = np.zeros((C, 8, 4), dtype=np.float32)
    ya = b.ProcessBlock(x[:T])
    wa = orc.biquad(x[:T], k, mode=MODE_FMA, state=ost)
    k2 = k.copy()
    k2[32:64] = k[::-1][32:64]
    b.SetCoeffs(k2[32:64], first=32)
    yb = b.ProcessBlock(x[T:])
    wb = orc.biquad(x[T:], k2, mode=MODE_FMA, state=ost)
    assert np.array_equal(bits(ya), bits(wa)) and np.array_equal(bits(yb), bits(wb))
    b.close()


def test_in_place_device_call_and_large_bank_properties(orc):
    """Full-width bank (65,536 channels): in-place device call; checks a sample of channels against the
    oracle and the identity / linearity properties on all of them."""
    C, T = 65536, 512
    dev = torch.empty((T, C), dtype=torch.float32, device="cuda")
    noise_device(dev.data_ptr(), 9, 0, C, 0, T)
    x = dev.cpu().numpy()
    k = c2_coefs(C)
    b = BiquadBank(C, 8)
    b.SetMode(MODE_FMA)
    b.SetCoeffs(k)
    b.ProcessDevice(dev.data_ptr(), dev.data_ptr(), T)  # in == out (main.cpp:82-85 calls Process in place)
    torch.cuda.synchronize()
    assert b.LastKernel() == "k_bq_tma"
    y = dev.cpu().numpy()
    pick = np.r_[0:64, 4000:4064, 65472:65536]
    want = orc.biquad(np.ascontiguousarray(x[:, pick]), np.ascontiguousarray(k[pick]), mode=MODE_FMA)
    assert np.array_equal(bits(y[:, pick]), bits(want))
    # channels c and c+64 share coefficients (c2_coefs is periodic in 64): same filter, different noise ->
    # outputs differ; identity coefficients pass the input through bit-exactly on every channel
    ident = np.zeros((C, 8, 5), dtype=np.float32)
    ident[:, :, 0] = 1.0
    b.Init(48000.0)
    b.SetCoeffs(ident)
    out = torch.empty_like(dev)
    noise_device(dev.data_ptr(), 9, 0, C, 0, T)
    b.ProcessDevice(dev.data_ptr(), out.data_ptr(), T)
    torch.cuda.synchronize()
    assert torch.equal(out, dev)
    b.close()


@pytest.mark.parametrize("S", [2, 4, 6, 8])
@pytest.mark.parametrize("mode", [MODE_EXACT, MODE_FMA])
@pytest.mark.parametrize("C", [132, 9600])
def test_systolic_shapes_bit_exact(orc, S, mode, C):
    """k_bq_sys: sections split over the warps of a CTA (scalar lanes below 9,472 channels, packed lanes
    above); ragged frame count, two calls so the per-warp history hand-over is exercised."""
    T = 777
    x = orc.noise(21, 0, C, 0, T)
    k = np.ascontiguousarray(c2_coefs(C)[:, :S])
    y, name, st = run_bank(x, k, mode, V_SYS, chunks=[500, 277])
    assert name == "k_bq_sys"
    ost = np.zeros((C, S, 4), dtype=np.float32)
    want = orc.biquad(x, k, mode=mode, state=ost)
    assert np.array_equal(bits(y), bits(want))
    assert np.array_equal(bits(st), bits(ost))


@pytest.mark.parametrize("S", [2, 4, 6, 8])
@pytest.mark.parametrize("mode", [MODE_EXACT, MODE_FMA, MODE_FAST])
@pytest.mark.parametrize("C,chunks", [(132, [500, 277]), (4096, [777]), (9600, [1, 7, 23, 24, 25, 697]), (36, [3, 774])])
def test_skewed_pairs_bit_exact(orc, S, mode, C, chunks):
    """k_bq_skew: two sections per warp in the halves of a packed FFMA2 stream, the second trailing the first; ragged
    frame counts (shorter than the lag, than a tile, straddling tiles) exercise the freezing path at both ends and the
    history hand-over between calls."""
    T = sum(chunks)
    x = orc.noise(22, 0, C, 0, T)
    k = np.ascontiguousarray(c2_coefs(C)[:, :S])
    y, name, st = run_bank(x, k, mode, V_SKEW, chunks=chunks)
    assert name == "k_bq_skew"
    ost = np.zeros((C, S, 4), dtype=np.float32)
    want = orc.biquad(x, k, mode=mode, state=ost)
    assert np.array_equal(bits(y), bits(want))
    assert np.array_equal(bits(st), bits(ost))


@pytest.mark.parametrize("C,chunks,cfg", [(4096, [3500, 3100, 3072], None), (4096, [4100, 4096, 4353], None), (8192, [1600, 2000, 1536], "96x8x4r"), (8192, [2048, 2300], "128x6x4r"), (1024, [3333], None),
                                          (4096, [700], "24x8x4r"), (4096, [1000], "48x8x4r"), (4096, [2000], "96x16x4r")])
def test_skewed_pairs_rolled_tile(orc, C, chunks, cfg, monkeypatch):
    """The FAST tier's rolled tile in every compiled height: 256 and 192 frames (<= 4 736 channels, blocks of 4 096 / 3 072 frames or more),
    128 and 96 (two CTAs per SM; since k_bq_roll took the 4 737 .. 9 472-channel banks these are reached by name), 48 and 24; several calls so that the pipeline's fill from and drain into the
    bank's history format is exercised at tile boundaries that move from call to call."""
    if cfg:
        monkeypatch.setenv("OLB_BQ_SKEW_CFG", cfg)
    T = sum(chunks)
    dev = torch.empty((T, C), dtype=torch.float32, device="cuda")
    noise_device(dev.data_ptr(), 24, 0, C, 0, T)
    out = torch.empty_like(dev)
    k = c2_coefs(C)
    b = BiquadBank(C, 8)
    b.SetMode(MODE_FAST)
    if cfg:
        b.SetVariant(V_SKEW)
    b.SetCoeffs(k)
    t0 = 0
    for n in chunks:
        b.ProcessDevice(dev[t0:].data_ptr(), out[t0:].data_ptr(), n)
        assert b.LastKernel() == "k_bq_skew"
        t0 += n
    torch.cuda.synchronize()
    pick = np.r_[0:34, C // 2 - 3:C // 2 + 30, C - 33:C]
    x = dev[:, torch.as_tensor(pick, device="cuda")].cpu().numpy()
    ost = np.zeros((len(pick), 8, 4), dtype=np.float32)
    want = orc.biquad(x, np.ascontiguousarray(k[pick]), mode=loader.MODE_FAST, state=ost)
    assert np.array_equal(bits(out.cpu().numpy()[:, pick]), bits(want))
    assert np.array_equal(bits(b.GetState()[pick]), bits(ost))
    b.close()


@pytest.mark.parametrize("C,chunks,cfg", [
    (8192, [1000, 700, 2049], "128x6x4c2"), (8192, [1000, 700, 2049], "128x6x5c2"), (8192, [1000, 700], "128x6x0c2"), (8192, [1000, 700], "128x6x1c2"),
    (8192, [300, 90], "32x6x4c2"), (8192, [513, 64, 1], "64x6x4c2"), (16384, [1100, 1000], "128x3x4c4"), (8192, [400], "32x6x5c4"), (8188, [900, 333], "96x8x4c2"),
    (4100, [500, 501], "64x6x4c4"), (8192, [1000, 700, 2049, 600, 100, 5000], None), (16384, [1100, 511, 1000], None), (6148, [3000], None)])
def test_channel_pair_kernel(orc, C, chunks, cfg, monkeypatch):
    """k_bq_roll (two channels per lane, 2 or 4 sections per warp chained in registers, rolled tile; stores in step or four
    positions late, one or two blocks per trip) in its compiled shapes: several calls -- the pipeline fills from and drains
    into the bank's history format at tile boundaries that move from call to call --, blocks shorter than a tile, a
    one-frame block, a channel count that leaves the last CTA partly empty.  Without a named shape the automatic choice
    takes it for blocks of 512 frames or more of 4 737 .. 9 472 and 14 209 .. 18 944-channel banks and k_bq_skew for the
    shorter ones: the two alternate on ONE bank, through the bank's common history format."""
    if cfg:
        monkeypatch.setenv("OLB_BQ_SKEW_CFG", cfg)
    T = sum(chunks)
    dev = torch.empty((T, C), dtype=torch.float32, device="cuda")
    noise_device(dev.data_ptr(), 25, 0, C, 0, T)
    out = torch.empty_like(dev)
    k = c2_coefs(C)
    b = BiquadBank(C, 8)
    b.SetMode(MODE_FAST)
    if cfg:
        b.SetVariant(V_SKEW)
    b.SetCoeffs(k)
    t0 = 0
    for n in chunks:
        b.ProcessDevice(dev[t0:].data_ptr(), out[t0:].data_ptr(), n)
        assert b.LastKernel() == ("k_bq_roll" if cfg or n >= 512 else "k_bq_skew"), (n, b.LastKernel())
        t0 += n
    torch.cuda.synchronize()
    pick = np.r_[0:66, C // 2 - 3:C // 2 + 30, C - 65:C]
    x = dev[:, torch.as_tensor(pick, device="cuda")].cpu().numpy()
    ost = np.zeros((len(pick), 8, 4), dtype=np.float32)
    want = orc.biquad(x, np.ascontiguousarray(k[pick]), mode=loader.MODE_FAST, state=ost)
    assert np.array_equal(bits(out.cpu().numpy()[:, pick]), bits(want))
    assert np.array_equal(bits(b.GetState()[pick]), bits(ost))
    b.close()


@pytest.mark.parametrize("mode", [MODE_EXACT, MODE_FMA])
@pytest.mark.parametrize("C,chunks", [(8192, [2400, 1536]), (4096, [4096, 3100]), (2048, [5000]), (16384, [2400])])
def test_mid_sized_banks_in_the_exact_and_fma_tiers_take_long_blocks(orc, mode, C, chunks):
    """The automatic choice for 512 .. 32 767-channel banks is mode-independent (k_bq_skew), its tile shape is not: the rolled
    tall tiles and k_bq_roll are the FAST tier's, EXACT and FMA run the unrolled 48-frame tiles whatever the block length."""
    T = sum(chunks)
    dev = torch.empty((T, C), dtype=torch.float32, device="cuda")
    noise_device(dev.data_ptr(), 26, 0, C, 0, T)
    out = torch.empty_like(dev)
    k = c2_coefs(C)
    b = BiquadBank(C, 8)
    b.SetMode(mode)
    b.SetCoeffs(k)
    t0 = 0
    for n in chunks:
        b.ProcessDevice(dev[t0:].data_ptr(), out[t0:].data_ptr(), n)
        assert b.LastKernel() == "k_bq_skew"
        t0 += n
    torch.cuda.synchronize()
    pick = np.r_[0:34, C // 2 - 3:C // 2 + 30, C - 33:C]
    x = dev[:, torch.as_tensor(pick, device="cuda")].cpu().numpy()
    ost = np.zeros((len(pick), 8, 4), dtype=np.float32)
    want = orc.biquad(x, np.ascontiguousarray(k[pick]), mode=mode, state=ost)
    assert np.array_equal(bits(out.cpu().numpy()[:, pick]), bits(want))
    assert np.array_equal(bits(b.GetState()[pick]), bits(ost))
    b.close()


@pytest.mark.parametrize("C", [8192, 16384])
def test_channel_pair_kernel_in_place(orc, C):
    """k_bq_roll with in == out (main.cpp:82-85 calls Process in place): a tile is stored 16 or 24 frames behind where it was
    loaded from, over rows the loader has already taken."""
    T = 3000
    dev = torch.empty((T, C), dtype=torch.float32, device="cuda")
    noise_device(dev.data_ptr(), 27, 0, C, 0, T)
    pick = np.r_[0:66, C // 2 - 3:C // 2 + 30, C - 65:C]
    x = dev[:, torch.as_tensor(pick, device="cuda")].cpu().numpy()
    k = c2_coefs(C)
    b = BiquadBank(C, 8)
    b.SetMode(MODE_FAST)
    b.SetCoeffs(k)
    b.ProcessDevice(dev.data_ptr(), dev.data_ptr(), 1700)
    b.ProcessDevice(dev[1700:].data_ptr(), dev[1700:].data_ptr(), T - 1700)
    torch.cuda.synchronize()
    assert b.LastKernel() == "k_bq_roll"
    want = orc.biquad(x, np.ascontiguousarray(k[pick]), mode=loader.MODE_FAST)
    assert np.array_equal(bits(dev.cpu().numpy()[:, pick]), bits(want))
    b.close()


def test_skewed_pairs_in_place(orc):
    C, T = 4096, 3000
    dev = torch.empty((T, C), dtype=torch.float32, device="cuda")
    noise_device(dev.data_ptr(), 23, 0, C, 0, T)
    x = dev.cpu().numpy()
    k = c2_coefs(C)
    b = BiquadBank(C, 8)
    b.SetMode(MODE_FAST)
    b.SetVariant(V_SKEW)
    b.SetCoeffs(k)
    b.ProcessDevice(dev.data_ptr(), dev.data_ptr(), T)
    torch.cuda.synchronize()
    assert b.LastKernel() == "k_bq_skew"
    pick = np.r_[0:40, 2000:2040, 4060:4096]
    want = orc.biquad(np.ascontiguousarray(x[:, pick]), np.ascontiguousarray(k[pick]), mode=loader.MODE_FAST)
    assert np.array_equal(bits(dev.cpu().numpy()[:, pick]), bits(want))
    b.close()


@pytest.mark.parametrize("variant,C,S", [(V_DIRECT, 77, 8), (V_TMA, 260, 8), (V_SYS, 132, 8), (V_SYS, 9600, 8), (V_TMA, 64, 3),
                                         (V_DIRECT, 64, 12), (V_SYS, 64, 4), (V_SKEW, 132, 8), (V_SKEW, 64, 4)])
def test_fast_mode_bit_exact_to_oracle_fast_and_close_to_exact(orc, variant, C, S):
    """OLB_MODE_FAST (normalised 4-op sections): same operations as the oracle's FAST mode -> bit-exact to it;
    against EXACT it is a tolerance tier (bound measured in DESIGN.md; both are compared with the float64 cascade)."""
    T = 1500
    x = orc.noise(31, 0, C, 0, T)
    k = np.ascontiguousarray(np.concatenate([c2_coefs(C), c2_coefs(C)[:, 4:]], axis=1)[:, :S])
    y, name, st = run_bank(x, k, MODE_FAST, variant, chunks=[700, 800])
    ost = np.zeros((C, S, 4), dtype=np.float32)
    want = orc.biquad(x, k, mode=loader.MODE_FAST, state=ost)
    assert np.array_equal(bits(y), bits(want)), name
    assert np.array_equal(bits(st), bits(ost))
    exact = orc.biquad(x, k, mode=MODE_EXACT)
    truth = orc.biquad_f64(x, k)
    # fp32 itself sits ~1e-4 from the float64 cascade on this coefficient set (30 Hz high-pass, pole radius 0.997):
    # FAST must be no further from the truth than EXACT is (x2 slack), and within 1e-3 of EXACT
    assert np.abs(y - truth).max() <= 2.0 * np.abs(exact - truth).max() + 1e-6
    assert np.abs(y - exact).max() <= 1e-3


def test_fast_mode_falls_back_to_fma_when_a_section_has_zero_b0(orc):
    C, T = 64, 400
    x = orc.noise(32, 0, C, 0, T)
    k = c2_coefs(C).copy()
    k[5, 3] = (0.0, 1.0, 0.0, 0.0, 0.0)  # a pure one-sample delay: b0 == 0 cannot be normalised
    y, _, _ = run_bank(x, k, MODE_FAST)
    assert np.array_equal(bits(y), bits(orc.biquad(x, k, mode=MODE_FMA)))


@pytest.mark.parametrize("variant,C", [(V_AUTO, 128), (V_TMA, 256), (V_SYS, 128), (V_SKEW, 128)])
def test_fast_history_follows_a_mid_stream_coefficient_change(orc, variant, C):
    """FAST keeps its history in the node scaling of the coefficient table; set_coeffs between two blocks (a filter sweep,
    an EQ gain change) re-expresses it at the next process call, as the oracle does (orc_biquad_rescale_state): bit-exact
    to the oracle across the change, and still a tolerance tier against EXACT (no jump)."""
    T = 600
    x = orc.noise(34, 0, C, 0, 2 * T)
    k = c2_coefs(C)
    k2 = k.copy()
    k2[32:96, 4] = k[::-1][32:96, 4]
    k2[32:96, 0, :3] *= 0.5
    b = BiquadBank(C, 8)
    b.SetMode(MODE_FAST)
    b.SetVariant(variant)
    b.SetCoeffs(k)
    ya = b.ProcessBlock(x[:T])
    b.SetCoeffs(k2[32:96], first=32)
    yb = b.ProcessBlock(x[T:])
    st = b.GetState()
    b.close()
    ost = np.zeros((C, 8, 4), dtype=np.float32)
    wa = orc.biquad(x[:T], k, mode=loader.MODE_FAST, state=ost)
    orc.biquad_rescale(ost, k, loader.MODE_FAST, k2, loader.MODE_FAST)
    wb = orc.biquad(x[T:], k2, mode=loader.MODE_FAST, state=ost)
    assert np.array_equal(bits(ya), bits(wa)) and np.array_equal(bits(yb), bits(wb))
    assert np.array_equal(bits(st), bits(ost))
    est = np.zeros((C, 8, 4), dtype=np.float32)
    orc.biquad(x[:T], k, mode=MODE_EXACT, state=est)
    assert np.abs(yb - orc.biquad(x[T:], k2, mode=MODE_EXACT, state=est)).max() <= 1e-3


def test_fast_fallback_is_not_sticky_and_mode_switches_rescale(orc):
    """A channel with b0 == 0 sends the bank to the FMA arithmetic; repairing it brings FAST back (history re-expressed in
    the table's scaling); switching the mode mid-stream goes through plain values again.  The oracle re-decides per call."""
    C, T = 64, 500
    x = orc.noise(35, 0, C, 0, 3 * T)
    k = c2_coefs(C)
    kbad = k.copy()
    kbad[5, 3] = (0.0, 1.0, 0.0, 0.0, 0.0)
    b = BiquadBank(C, 8)
    b.SetMode(MODE_FAST)
    b.SetCoeffs(kbad)
    ost = np.zeros((C, 8, 4), dtype=np.float32)
    y1 = b.ProcessBlock(x[:T])
    w1 = orc.biquad(x[:T], kbad, mode=loader.MODE_FAST, state=ost)          # runs FMA: channel 5 cannot be normalised
    b.SetCoeffs(k[5:6], first=5)
    y2 = b.ProcessBlock(x[T:2 * T])
    orc.biquad_rescale(ost, kbad, loader.MODE_FAST, k, loader.MODE_FAST)
    w2 = orc.biquad(x[T:2 * T], k, mode=loader.MODE_FAST, state=ost)
    b.SetMode(MODE_EXACT)
    y3 = b.ProcessBlock(x[2 * T:])
    orc.biquad_rescale(ost, k, loader.MODE_FAST, k, loader.MODE_EXACT)
    w3 = orc.biquad(x[2 * T:], k, mode=MODE_EXACT, state=ost)
    b.close()
    assert np.array_equal(bits(y1), bits(w1)) and np.array_equal(bits(y1), bits(orc.biquad(x[:T], kbad, mode=MODE_FMA)))
    assert np.array_equal(bits(y2), bits(w2))
    assert np.array_equal(bits(y3), bits(w3))


def test_control_calls_are_ordered_on_a_non_blocking_stream(orc):
    """process on a non-default (non-blocking) stream, Init, process again: Init takes effect between the two blocks and
    cannot be overtaken by the first block's final history write; a coefficient change reaches only the second block."""
    C, T = 9600, 4000
    s = torch.cuda.Stream()
    dev = torch.empty((T, C), dtype=torch.float32, device="cuda")
    out1, out2, out3 = torch.empty_like(dev), torch.empty_like(dev), torch.empty_like(dev)
    noise_device(dev.data_ptr(), 36, 0, C, 0, T)
    torch.cuda.synchronize()
    k = c2_coefs(C)
    b = BiquadBank(C, 8)
    b.SetMode(MODE_FMA)
    b.SetCoeffs(k)
    b.ProcessDevice(dev.data_ptr(), out1.data_ptr(), T, LAYOUT_FRAMES, s.cuda_stream)
    b.Init(48000.0)
    b.ProcessDevice(dev.data_ptr(), out2.data_ptr(), T, LAYOUT_FRAMES, s.cuda_stream)
    b.SetCoeffs(k[::-1].copy())
    b.ProcessDevice(dev.data_ptr(), out3.data_ptr(), T, LAYOUT_FRAMES, torch.cuda.current_stream().cuda_stream)  # another stream
    torch.cuda.synchronize()
    assert torch.equal(out1, out2)
    x = dev[:, :64].cpu().numpy()
    st = np.zeros((64, 8, 4), dtype=np.float32)
    orc.biquad(x, k[:64], mode=MODE_FMA, state=st)
    want3 = orc.biquad(x, k[::-1][:64], mode=MODE_FMA, state=st)
    assert np.array_equal(bits(out3[:, :64].cpu().numpy()), bits(want3))
    b.close()


def test_fast_mode_one_pole_kinds(orc):
    C, T = 64, 600
    x = orc.noise(33, 0, C, 0, T)
    k = c2_coefs(C)[:, :4].copy()
    k[:, 1, 0] = 0.85
    k[:, 3, 0] = 0.999
    kinds = np.array([0, KIND_LPF, 0, KIND_PORT], dtype=np.uint8)
    y, _, _ = run_bank(x, k, MODE_FAST, kinds=kinds)
    assert np.array_equal(bits(y), bits(orc.biquad(x, k, kinds=kinds, mode=loader.MODE_FAST)))
    assert np.abs(y - orc.biquad(x, k, kinds=kinds, mode=MODE_EXACT)).max() <= 1e-3


@pytest.mark.parametrize("mode", [MODE_EXACT, MODE_FAST])
def test_cpp_host_mirror_matches_oracle(orc, mode, tmp_path):
    """A C++ caller written against include/ol_gpu/BiquadBank.hpp (Init -> SetCoeffs -> Process per frame -> ProcessBlock)."""
    import os
    import subprocess

    import __graft_entry__ as g
    g.build_cpp_hosts()
    C, T = 256, 500
    x = orc.noise(41, 0, C, 0, T)
    k = c2_coefs(C)
    x.tofile(tmp_path / "in.f32")
    k.tofile(tmp_path / "k.f32")
    exe = os.path.join(g.ROOT, "tests", "cpp", "biquad_host")
    r = subprocess.run([exe, str(C), str(T), str(tmp_path / "in.f32"), str(tmp_path / "k.f32"), str(tmp_path / "out.f32"), str(mode)],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    y = np.fromfile(tmp_path / "out.f32", dtype=np.float32).reshape(T, C)
    assert np.array_equal(bits(y), bits(orc.biquad(x, k, mode=mode)))


def test_errors():
    b = BiquadBank(8, 2)
    with pytest.raises(OlbError):
        b.SetCoeffs(np.zeros((9, 2, 5), dtype=np.float32))
    with pytest.raises(OlbError):
        b.SetKinds([0, 9])
    with pytest.raises(OlbError):
        b.SetVariant(V_TMA) or b.SetKinds([1, 0]) or b.ProcessBlock(np.zeros((4, 8), dtype=np.float32))
    b.close()
    assert _lib.launch_count() > 0
