"""CPU: known-answer tests that pin oracle/h264_core.c to ITU-T H.264 (the x264 half of the oracle has no
reference vectors to lean on -- the reference repository has no tests and libx264 is absent -- so the pins are
identities and tables taken from the standard)."""
import numpy as np
import pytest

import oracle_lib as O

RNG = np.random.Generator(np.random.PCG64(42))


def test_levelscale_and_quant_tables_follow_the_standard():
    # H.264 8.5.9: normAdjust4x4 rows m=0..5: (10,16,13) ... ; LevelScale = 16 * normAdjust (flat matrix)
    v = [(10, 16, 13), (11, 18, 14), (13, 20, 16), (14, 23, 18), (16, 25, 20), (18, 29, 23)]
    lib = O.load()
    for m in range(6):
        assert lib.gvc_oracle_dequant4_scale(m, 0) == 16 * v[m][0]        # (0,0)
        assert lib.gvc_oracle_dequant4_scale(m + 12, 5) == 16 * v[m][1]   # (1,1)
        assert lib.gvc_oracle_dequant4_scale(m + 6, 1) == 16 * v[m][2]    # (0,1)
    # forward multipliers: A(qp%6) * LevelScale ~= 2^15+6 / normalisation (a^2=1/4 -> 1, ab/2, b^2/4)
    for m in range(6):
        mf, _ = O.quant4_tables(m, False)       # qp < 6: mf = 2 * table value
        assert abs(int(mf[0]) // 2 * v[m][0] - (1 << 17)) < v[m][0]
        assert abs(int(mf[5]) // 2 * v[m][1] - (1 << 17) * 16 / 25) < v[m][1]
        assert abs(int(mf[1]) // 2 * v[m][2] - (1 << 17) * 4 / 5) < v[m][2]
    # every +6 in qp halves the multiplier (rounded)
    for qp in range(6, 52):
        a, _ = O.quant4_tables(qp, True)
        b, _ = O.quant4_tables(qp - 6, True)
        assert np.all(np.abs(a.astype(int) * 2 - b.astype(int)) <= 1)
    # dead zones: intra rounds at 1/3, inter at 1/6 (x264: 21/64 and 11/64)
    mf, bi = O.quant4_tables(24, True)
    assert abs(int(bi[0]) * int(mf[0]) / 65536 - 21 / 64) < 0.01
    mf, bi = O.quant4_tables(24, False)
    assert abs(int(bi[0]) * int(mf[0]) / 65536 - 11 / 64) < 0.01
    # chroma QP table (H.264 Table 8-15)
    assert [lib.gvc_oracle_chroma_qp(q) for q in (0, 29, 30, 34, 39, 45, 51)] == [0, 29, 29, 32, 35, 38, 39]


def test_dct4_flat_and_basis():
    pred = np.zeros((4, 4), np.uint8)
    c = O.sub_dct(np.full((4, 4), 9, np.uint8), pred, 4)
    assert c[0, 0] == 16 * 9 and np.count_nonzero(c) == 1
    # Cf * X * Cf^T with Cf rows (1,1,1,1),(2,1,-1,-2),(1,-1,-1,1),(1,-2,2,-1)  (H.264 8.5.12 inverse of)
    Cf = np.array([[1, 1, 1, 1], [2, 1, -1, -2], [1, -1, -1, 1], [1, -2, 2, -1]])
    for _ in range(50):
        s = RNG.integers(0, 256, (4, 4), dtype=np.uint8)
        p = RNG.integers(0, 256, (4, 4), dtype=np.uint8)
        d = s.astype(int) - p.astype(int)
        np.testing.assert_array_equal(O.sub_dct(s, p, 4), Cf @ d @ Cf.T)


def test_idct4_matches_spec_matrix_form_on_dequantised_multiples():
    # for coefficients that are multiples of 64 the >>1 steps are exact, so the integer butterflies must
    # equal Ci * W * Ci^T / 64 with Ci = [[1,1,1,.5],[1,.5,-1,-1],[1,-.5,-1,1],[1,-1,1,-.5]]
    Ci = np.array([[1, 1, 1, .5], [1, .5, -1, -1], [1, -.5, -1, 1], [1, -1, 1, -.5]])
    pred = np.full((4, 4), 128, np.uint8)
    for _ in range(50):
        w = RNG.integers(-6, 7, (4, 4)) * 64
        want = np.clip(128 + np.floor((Ci @ w @ Ci.T + 32) / 64), 0, 255).astype(np.uint8)
        np.testing.assert_array_equal(O.add_idct(pred, w, 4), want)


@pytest.mark.parametrize("n", [4, 8])
def test_transform_quant_roundtrip_is_near_lossless_at_qp0_and_bounded_elsewhere(n):
    for qp in (0, 12, 23, 36, 51):
        mf, bias = (O.quant4_tables if n == 4 else O.quant8_tables)(qp, True)
        worst = 0
        for _ in range(40):
            s = RNG.integers(0, 256, (n, n), dtype=np.uint8)
            p = RNG.integers(0, 256, (n, n), dtype=np.uint8)
            c = O.sub_dct(s, p, n)
            q, _ = O.quant(c, mf, bias)
            r = O.add_idct(p, O.dequant(q.reshape(n, n), qp, n), n)
            worst = max(worst, int(np.abs(r.astype(int) - s.astype(int)).max()))
        qstep = 0.625 * 2 ** (qp / 6)
        assert worst <= max(1, 2.5 * qstep), (qp, worst)
        if qp == 0:
            assert worst <= 1


def test_dc_hadamards():
    x = RNG.integers(-2000, 2000, 16).astype(np.int16) * 2      # even values: the (x+1)>>1 rounding is exact
    y = O.hadamard_dc(x, "dct4")
    H = np.array([[1, 1, 1, 1], [1, 1, -1, -1], [1, -1, -1, 1], [1, -1, 1, -1]])
    np.testing.assert_array_equal(y.reshape(4, 4), (H @ x.reshape(4, 4).astype(int) @ H.T) // 2)
    small = RNG.integers(-100, 100, 16).astype(np.int16)
    np.testing.assert_array_equal(O.hadamard_dc(O.hadamard_dc(small * 2, "dct4"), "idct4"), small.astype(int) * 16)
    d = np.array([5, -3, 7, 11], np.int16)
    np.testing.assert_array_equal(O.hadamard_dc(O.hadamard_dc(d, "dct2"), "dct2"), d * 4)
    assert list(O.hadamard_dc(d, "dct2")) == [20, 4, -16, 12]     # [[a+b+c+d, a-b+c-d],[a+b-c-d, a-b-c+d]]


def test_sad_satd_identities():
    a = RNG.integers(0, 256, (16, 16), dtype=np.uint8)
    assert O.sad(a, a) == 0 and O.satd(a, a) == 0
    b = np.clip(a.astype(int) + 3, 0, 255).astype(np.uint8)
    a2 = np.clip(a, 0, 252).astype(np.uint8)
    b2 = (a2 + 3).astype(np.uint8)
    assert O.sad(a2, b2) == 3 * 256
    assert O.satd(a2, b2) == 16 * (16 * 3 // 2)          # per 4x4: only the DC term, 16*3, halved
    # SATD of a 4x4 block equals sum |H D H| / 2
    H = np.array([[1, 1, 1, 1], [1, 1, -1, -1], [1, -1, -1, 1], [1, -1, 1, -1]])
    for _ in range(20):
        s, p = RNG.integers(0, 256, (4, 4), dtype=np.uint8), RNG.integers(0, 256, (4, 4), dtype=np.uint8)
        d = s.astype(int) - p.astype(int)
        assert O.satd(s, p) == np.abs(H @ d @ H.T).sum() // 2
    # planted shift: SAD is minimal exactly at the planted displacement
    ref = RNG.integers(0, 256, (48, 48), dtype=np.uint8)
    cur = ref[16 + 3:32 + 3, 16 - 5:32 - 5]
    costs = {(dx, dy): O.sad(cur, ref[16 + dy:32 + dy, 16 + dx:32 + dx]) for dx in range(-8, 9) for dy in range(-8, 9)}
    assert min(costs, key=costs.get) == (-5, 3) and costs[(-5, 3)] == 0


def test_hpel_filter_known_answers():
    # constant plane stays constant; ramp p=x gives H = x+1 (6-tap on a linear ramp = exact midpoint, rounded up)
    flat = np.full((24, 40), 77, np.uint8)
    for p in O.hpel_planes(flat, 32)[1:]:
        assert (p[29:-29, 29:-29] == 77).all()
    ramp = np.tile(np.arange(40, dtype=np.uint8) * 4, (24, 1))
    full, H, V, C = O.hpel_planes(ramp, 32)
    np.testing.assert_array_equal(H[32:56, 32 + 3:32 + 36], ramp[:, 3:36] + 2)
    np.testing.assert_array_equal(V[32:56, 32:72], ramp)
    # impulse of 32 -> H plane shows the tap vector (1,-5,20,20,-5,1) clipped at 0
    imp = np.zeros((16, 32), np.uint8)
    imp[8, 16] = 32
    full, H, V, C = O.hpel_planes(imp, 32)
    assert list(H[32 + 8, 32 + 13:32 + 19]) == [1, 0, 20, 20, 0, 1]
    assert list(V[32 + 5:32 + 11, 32 + 16]) == [1, 0, 20, 20, 0, 1]
    # centre plane of an impulse of 255: outer product of the tap vector, (x*255 + 512) >> 10, clipped
    imp[8, 16] = 255
    full, H, V, C = O.hpel_planes(imp, 32)
    t = np.array([1, -5, 20, 20, -5, 1])
    want = np.clip((np.outer(t, t) * 255 + 512) >> 10, 0, 255)
    np.testing.assert_array_equal(C[32 + 5:32 + 11, 32 + 13:32 + 19], want)


def test_lowres_and_mc_known_answers():
    flat = np.full((32, 48), 200, np.uint8)
    for p in O.lowres_planes(flat):
        assert (p == 200).all()
    img = RNG.integers(0, 256, (32, 48), dtype=np.uint8)
    l0 = O.lowres_planes(img)[0]
    a, b, c, d = (img[0::2, 0::2].astype(int), img[1::2, 0::2].astype(int), img[0::2, 1::2].astype(int), img[1::2, 1::2].astype(int))
    np.testing.assert_array_equal(l0, (((a + b + 1) >> 1) + ((c + d + 1) >> 1) + 1) >> 1)
    planes = O.hpel_planes(img, 32)
    full, H, V, C = planes
    blk = lambda P, x, y: P[32 + y:32 + y + 8, 32 + x:32 + x + 8].astype(int)
    np.testing.assert_array_equal(O.mc_luma(planes, 32, 8, 8, 4 * 3, -4 * 2, 8, 8), blk(full, 11, 6))       # integer mv
    np.testing.assert_array_equal(O.mc_luma(planes, 32, 8, 8, 2, 0, 8, 8), blk(H, 8, 8))                    # b
    np.testing.assert_array_equal(O.mc_luma(planes, 32, 8, 8, 0, 2, 8, 8), blk(V, 8, 8))                    # h
    np.testing.assert_array_equal(O.mc_luma(planes, 32, 8, 8, 2, 2, 8, 8), blk(C, 8, 8))                    # j
    np.testing.assert_array_equal(O.mc_luma(planes, 32, 8, 8, 1, 0, 8, 8), (blk(full, 8, 8) + blk(H, 8, 8) + 1) >> 1)   # a
    np.testing.assert_array_equal(O.mc_luma(planes, 32, 8, 8, 3, 0, 8, 8), (blk(full, 9, 8) + blk(H, 8, 8) + 1) >> 1)   # c
    np.testing.assert_array_equal(O.mc_luma(planes, 32, 8, 8, 1, 1, 8, 8), (blk(H, 8, 8) + blk(V, 8, 8) + 1) >> 1)      # e
    np.testing.assert_array_equal(O.mc_luma(planes, 32, 8, 8, 3, 3, 8, 8), (blk(H, 8, 9) + blk(V, 9, 8) + 1) >> 1)      # r
    np.testing.assert_array_equal(O.mc_luma(planes, 32, 8, 8, 2, 3, 8, 8), (blk(C, 8, 8) + blk(H, 8, 9) + 1) >> 1)      # q
    ch = O.pad_plane(RNG.integers(0, 256, (16, 24), dtype=np.uint8), 16)
    cb = lambda x, y: ch[16 + y:16 + y + 4, 16 + x:16 + x + 4].astype(int)
    np.testing.assert_array_equal(O.mc_chroma(ch, 16, 4, 4, 8, -16, 4, 4), cb(5, 2))                         # whole chroma samples
    np.testing.assert_array_equal(O.mc_chroma(ch, 16, 4, 4, 4, 0, 4, 4), (cb(4, 4) + cb(5, 4) + 1) >> 1)     # half way
    np.testing.assert_array_equal(O.mc_chroma(ch, 16, 4, 4, 2, 6, 4, 4),
                                  (12 * cb(4, 4) + 4 * cb(5, 4) + 36 * cb(4, 5) + 12 * cb(5, 5) + 32) >> 6)


def test_nal_wrapper_escapes_exactly_like_the_standard_state_machine():
    """H.264 7.4.1 / B.1: 0x03 goes in front of any byte <= 3 that follows two zero bytes.  The product's wrapper skips from zero
    byte to zero byte; compare it with the byte-by-byte state machine on payloads full of zeros and on random ones."""
    import emu_lib
    rng = np.random.Generator(np.random.PCG64(7))

    def reference(rbsp, ref_idc, t):
        out, zeros = bytearray([0, 0, 0, 1, (ref_idc << 5) | t]), 0
        for b in rbsp:
            if zeros >= 2 and b <= 3:
                out.append(3)
                zeros = 0
            out.append(b)
            zeros = zeros + 1 if b == 0 else 0
        return bytes(out)

    cases = [b"", b"\x00", b"\x00\x00", b"\x00\x00\x00", b"\x00\x00\x01", b"\x00\x00\x03\x00\x00\x03", b"\x00" * 17, b"\x01\x00\x00\x04\x00\x00\x02",
             bytes(range(256)) * 3]
    for n, zero_share in ((1, 0.5), (7, 0.9), (300, 0.6), (5000, 0.3), (5000, 0.02), (100000, 0.004)):
        for _ in range(4):
            v = rng.integers(0, 256, n, dtype=np.uint8)
            v[rng.random(n) < zero_share] = 0
            v[rng.random(n) < zero_share / 4] = rng.integers(1, 4)
            cases.append(v.tobytes())
    for c in cases:
        assert emu_lib.append_nal(c, 2, 1) == reference(c, 2, 1), "payload of %d bytes" % len(c)
