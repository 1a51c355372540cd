"""CPU tests: the oracle (oracle/vo_oracle.c) against the committed golden vectors, against Python cv2 for the
third-party blur, and - where oracle/_ref/libvoref.so exists - against the reference's own code live."""
import numpy as np
import pytest

from conftest import dense_R, unpack

KAT_SETS = ("random", "ties", "n2_is_1", "n2_is_2", "near_duplicates", "noisy_duplicates", "prefix_ties")


def test_blur_matches_cv2(oracle, factory):
    """cv::GaussianBlur(u8, 7x7, 0) == (sum w w src + 2048) >> 12 with REFLECT_101; cv2 version recorded."""
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(3)
    imgs = [factory[0], rng.integers(0, 256, (376, 1241), dtype=np.uint8), rng.integers(0, 256, (97, 133), dtype=np.uint8),
            rng.integers(0, 256, (9, 8), dtype=np.uint8), np.full((40, 50), 255, np.uint8)]
    for img in imgs:
        assert np.array_equal(oracle.blur(img), cv2.GaussianBlur(img, (7, 7), 0)), f"cv2 {cv2.__version__}"


@pytest.mark.parametrize("tag,idx", [("factory1", 0), ("factory2", 1)])
def test_oracle_stages_vs_golden_factory(oracle, factory, tag, idx):
    img, g = factory[idx], factory[2]
    kp, desc, ang, blurred, R = oracle.extract(img, 25000.0, 5, 1500, stages=True)
    assert np.array_equal(blurred, g[f"{tag}_blurred"])
    assert np.array_equal(R, dense_R(g[f"{tag}_R_idx"], g[f"{tag}_R_val"], img.shape))  # bit-exact fp32
    assert np.array_equal(kp, g[f"{tag}_kp"].astype(np.float32))
    assert np.array_equal(oracle.nms(R, 3, 2000), g[f"{tag}_kp_nms3_2000"].astype(np.float32))
    # orientation/descriptor go through libm atan2f/sinf/cosf: identical on the machine that made the fixtures,
    # allowed to differ by an ulp (and a handful of bits) on another libm build
    assert np.allclose(ang, g[f"{tag}_angles"], rtol=0, atol=1e-6)
    assert (desc != unpack(g[f"{tag}_desc"])).mean() < 1e-4


def test_oracle_pipeline_vs_golden_kitti(oracle, kitti_pair):
    fr, g = kitti_pair
    kp0, d0, ang, blurred, R = oracle.extract(fr[0], 25000.0, 5, 1500, stages=True)
    kp1, d1 = oracle.extract(fr[1], 25000.0, 5, 1500)
    assert np.array_equal(blurred, g["kitti0_blurred"])
    assert np.array_equal(R, dense_R(g["kitti0_R_idx"], g["kitti0_R_val"], fr[0].shape))
    assert np.array_equal(kp0, g["kitti0_kp"].astype(np.float32)) and np.array_equal(kp1, g["kitti1_kp"].astype(np.float32))
    assert (d0 != unpack(g["kitti0_desc"])).mean() < 1e-4
    m = oracle.match(unpack(g["kitti0_desc"]), unpack(g["kitti1_desc"]), 32)
    assert np.array_equal(m, g["kitti_matches_ref32"])


def test_oracle_matcher_vs_golden(oracle, factory, matcher_kat):
    g = factory[2]
    d1, d2 = unpack(g["factory1_desc"]), unpack(g["factory2_desc"])
    assert np.array_equal(oracle.match(d1, d2, 32), g["factory_matches_ref32"])
    assert len(g["factory_matches_ref32"]) == 559 and g["factory_matches_ref32"][0].tolist() == [5, 924]
    assert np.array_equal(oracle.match(d1, d2, 512), g["factory_matches_full512_ref"])
    for name in KAT_SETS:
        a, b = unpack(matcher_kat[f"{name}_d1"]), unpack(matcher_kat[f"{name}_d2"])
        assert np.array_equal(oracle.match(a, b, 32), matcher_kat[f"{name}_ref32"]), name
        assert np.array_equal(oracle.match(a, b, 512), matcher_kat[f"{name}_full512_ref"]), name


def test_matcher_semantics(oracle):
    """Appendix A7: first index of the minimum wins; d2 is the second smallest of the multiset; N2 < 2 => no match."""
    z = np.zeros((1, 512), np.uint8)
    t = np.zeros((3, 512), np.uint8)
    t[0, :4] = 1; t[1, :4] = 1; t[2, :20] = 1          # distances 4, 4, 20 -> tie at the minimum: d2 = d1 -> rejected
    pairs, raw = oracle.match(z, t, 32, raw=True)
    assert len(pairs) == 0 and raw[0].tolist() == [4, 0, 4]
    t[1, :12] = 1                                       # 4, 12, 20 -> 4 < 0.75*12
    pairs, raw = oracle.match(z, t, 32, raw=True)
    assert pairs.tolist() == [[0, 0]] and raw[0].tolist() == [4, 0, 12]
    t2 = np.zeros((2, 512), np.uint8); t2[1, 100:140] = 1   # tests >= 32 are invisible to REF32, visible to FULL512
    assert len(oracle.match(z, t2, 32)) == 0 and oracle.match(z, t2, 512).tolist() == [[0, 0]]
    assert len(oracle.match(z, t[:1], 32)) == 0 and len(oracle.match(z, t[:0], 32)) == 0 and len(oracle.match(t[:0], t, 32)) == 0


def test_nms_semantics(oracle):
    """Appendix A4: strict maximum (plateaus rejected), margins 35/37, top-N by (R, i, j) descending, (y, x) order."""
    R = np.zeros((120, 140), np.float32)
    R[50, 60] = 5.0; R[50, 61] = 5.0            # plateau: both rejected
    R[60, 37] = 3.0; R[60, 36] = 2.0            # j = 37 is inside the margin, j = 36 is not (and is a smaller neighbour)
    R[35, 80] = 4.0; R[34, 90] = 9.0            # i = 35 allowed, i = 34 not
    R[85, 103] = 7.0; R[86, 100] = 1.0          # i <= H-35 = 85, j <= W-37 = 103
    R[70, 70] = 6.0; R[70, 73] = 6.0; R[73, 70] = 6.0   # equal responses, far enough apart for k = 5
    kp = oracle.nms(R, 5, 100)
    assert kp.tolist() == [[80, 35], [37, 60], [70, 70], [73, 70], [70, 73], [103, 85]]
    top2 = oracle.nms(R, 5, 2)                  # 7.0 first, then among the 6.0 ties the largest (i, j) = (73, 70)
    assert top2.tolist() == [[70, 73], [103, 85]]
    top3 = oracle.nms(R, 5, 3)                  # next tie: (70, 73) beats (70, 70)
    assert top3.tolist() == [[73, 70], [70, 73], [103, 85]]
    assert len(oracle.nms(R, 3, 100)) == 6 and len(oracle.nms(np.zeros((80, 90), np.float32), 5, 10)) == 0
    assert len(oracle.nms(np.ones((30, 40), np.float32), 5, 10)) == 0   # smaller than the margins


def test_response_borders_and_quirks(oracle):
    rng = np.random.default_rng(5)
    b = rng.integers(0, 256, (40, 50), dtype=np.uint8)
    jx, jy, jxy = oracle.gradients(b)
    for j in (jx, jy, jxy):
        assert not j[0].any() and not j[-1].any() and not j[:, 0].any() and not j[:, -1].any()
    i, k = 7, 9
    sx = [int(b[i - 1, k + d]) - int(b[i + 1, k + d]) for d in (-1, 0, 1)]
    sy = [int(b[i - 1, k + d]) + 2 * int(b[i, k + d]) + int(b[i + 1, k + d]) for d in (-1, 0, 1)]
    assert jx[i, k] == sx[0] + 2 * sx[1] + sx[2] and jy[i, k] == sy[0] - sy[2] and jxy[i, k] == sx[0] - sx[2]
    R = oracle.response(b, 0.0)
    assert not R[:2].any() and not R[-2:].any() and not R[:, :2].any() and not R[:, -2:].any()
    a = float((jx[5:10, 7:12] ** 2).sum()); bb = float((jy[5:10, 7:12] ** 2).sum()); c = float(jxy[5:10, 7:12].sum())
    f = np.float32
    det = f(f(f(a) * f(bb)) - f(f(c) * f(c))); tr = f(f(a) + f(bb))
    disc = f(f(tr * tr) - f(f(4) * det))
    want = f(f(tr / f(2)) - f(f(0.5) * np.sqrt(disc))) if disc >= 0 else f(0)
    want = want if want > 0 else f(0)
    assert R[7, 9] == want


def test_restatement_equals_reference_live(oracle, reference, factory, kitti_pair):
    """Pin: oracle/vo_oracle.c == the reference's own translation units, stage by stage, on fresh inputs."""
    rng = np.random.default_rng(17)
    noise = rng.integers(0, 256, (150, 170), dtype=np.uint8)
    for img in (factory[0], kitti_pair[0][1], noise):
        bgr = np.repeat(img[:, :, None], 3, axis=2)
        assert np.array_equal(reference.bgr2gray(bgr), img) and np.array_equal(oracle.bgr2gray(bgr), img)
        blurred = reference.blur(img)
        assert np.array_equal(blurred, oracle.blur(img))
        for a, b in zip(reference.gradients(blurred), oracle.gradients(blurred)):
            assert np.array_equal(a, b)
        R = reference.response(blurred)
        assert np.array_equal(R, oracle.response(blurred, 25000.0))
        for k, n in ((5, 1500), (3, 2000), (5, 100), (7, 50)):
            assert np.array_equal(reference.nms(R, k, n), oracle.nms(R, k, n))
        kp_r, d_r = reference.extract(img, threads=4)
        kp_o, d_o = oracle.extract(img, 25000.0, 5, 1500)
        assert np.array_equal(kp_r, kp_o) and np.array_equal(d_r, d_o)
    colour = rng.integers(0, 256, (60, 70, 3), dtype=np.uint8)
    assert np.array_equal(reference.bgr2gray(colour), oracle.bgr2gray(colour))
    d1 = rng.integers(0, 2, (130, 512), dtype=np.uint8); d2 = rng.integers(0, 2, (90, 512), dtype=np.uint8)
    d2[:40] = d1[:40]; d2[:40, rng.integers(0, 32, 40)] ^= 1
    for thr in (1, 3, 8):
        assert np.array_equal(reference.match(d1, d2, threads=thr), oracle.match(d1, d2, 32))


def test_full512_restatement_equals_reference_serial_live(oracle, reference_serial, factory, matcher_kat):
    """Pin of the FULL512 mode: oracle/vo_oracle.c against matching_serial.cpp:24-77 compiled in place
    (hammingDistance over all 512 bytes, matchCustomBinaryDescriptorsSerial), on the KATs, the factory pair and fresh sets."""
    g = factory[2]
    rng = np.random.default_rng(29)
    a = rng.integers(0, 2, (150, 512), dtype=np.uint8)
    b = a[rng.permutation(150)][:120].copy()
    b[rng.random(b.shape) < 0.3] ^= 1                      # distances on both sides of the 0.75 ratio
    b[:10] = b[10:20]                                       # exact duplicates in the train set: d2 == d1
    sets = [(unpack(g["factory1_desc"]), unpack(g["factory2_desc"])), (a, b), (a[:1], b), (a, b[:2]), (a, b[:1])]
    sets += [(unpack(matcher_kat[f"{n}_d1"]), unpack(matcher_kat[f"{n}_d2"])) for n in KAT_SETS]
    for d1, d2 in sets:
        for ratio in (0.75, 0.9, 0.5):
            assert np.array_equal(reference_serial.match(d1, d2, ratio), oracle.match(d1, d2, 512, ratio))
    _, raw = oracle.match(a, b, 512, raw=True)
    for q in (0, 7, 149):                                   # best distance of the restatement == the reference's hammingDistance
        assert raw[q, 0] == reference_serial.hamming(a[q], b[raw[q, 1]])
        assert raw[q, 0] == min(reference_serial.hamming(a[q], row) for row in b)


def test_reference_strip_race_is_avoided_by_single_strip(reference, factory):
    """Single strip, any thread count => identical output (the parity setting)."""
    a = reference.extract(factory[0], threads=1)
    b = reference.extract(factory[0], threads=8)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])


def test_reciprocal_division_is_exact(oracle):
    """The FREAK kernel replaces (ic*d)/|d| by q=m*rn; r=fma(-q,|d|,m); q'=fma(r,rn,q).  Exhaustive proof over the
    whole input domain (903 pairs x 511 intensity differences x 2 axes) that q' is the IEEE quotient."""
    bad, total = oracle.check_recip_division()
    assert total == 903 * 511 * 2 and bad == 0


def test_split_weight_product_is_exact(oracle):
    """The FREAK kernel evaluates RN((ic*d)/|d|) as fma(ic, wh, RN(ic*wl)), (wh, wl) the two-float split of d/|d|.
    Exhaustive proof over the whole input domain (903 pairs x 511 intensity differences x 2 axes)."""
    bad, total = oracle.check_split_weight()
    assert total == 903 * 511 * 2 and bad == 0



def test_heap_order_rule_equals_the_references_own_output_order(oracle, reference, factory):
    """The reference's non_maximum_suppression returns keypoints as its max-heap of (R, i, j) pops them.  The adapter's
    CornerDetectionParallel::non_maximum_suppression facade restores that order from the (y, x)-sorted GPU list; the
    rule it uses (tests/conftest.py::heap_order) is pinned here to the reference's own code, on the factory response map
    and on a map with equal responses (ties go to the larger row, then the larger column)."""
    from conftest import heap_order
    if not hasattr(reference.lib, "voref_nms_ranked"):
        pytest.skip("oracle/_ref/libvoref.so predates voref_nms_ranked: rebuild it with /root/reference present")
    R = oracle.response(oracle.blur(factory[0]), 25000.0)
    for window, n in ((5, 1500), (3, 2000), (5, 40)):
        ranked = reference.nms_ranked(R, window, n)
        assert len(ranked) == len(reference.nms(R, window, n)) > 0
        assert np.array_equal(ranked, heap_order(reference.nms(R, window, n), R)), (window, n)
        assert np.array_equal(ranked, heap_order(oracle.nms(R, window, n), R)), (window, n)
        assert (np.diff(R[ranked[:, 1].astype(int), ranked[:, 0].astype(int)]) <= 0).all()      # strongest first
    T = np.zeros((120, 160), np.float32)
    for (i, j) in ((40, 50), (40, 90), (70, 50), (70, 120), (50, 70)):
        T[i, j] = 30000.0
    T[60, 100] = 40000.0
    ranked = reference.nms_ranked(T, 5, 10)
    assert ranked.tolist() == [[100, 60], [120, 70], [50, 70], [70, 50], [90, 40], [50, 40]]
    assert np.array_equal(ranked, heap_order(oracle.nms(T, 5, 10), T))
