"""GPU parity: preprocess::txt2pc (src/smoothing.cpp:39-81) parsed on the device — bit-identical to the reference's
own stream statements run by the C++ library of this image (oracle/textio.cpp)."""
import numpy as np
import pytest

from pointcloudprocessing_b200 import synth

pytestmark = pytest.mark.gpu


def _check(pcp, oracle, text: bytes):
    ref = oracle.txt2pc(text)
    got = pcp.set_cloud_txt(text)
    assert got.shape == ref.shape
    bad = np.nonzero((got.view(np.uint32) != ref.view(np.uint32)).any(1))[0]
    assert bad.size == 0, f"{bad.size} lines differ, first {bad[:5]}: got {got[bad[:3]]} ref {ref[bad[:3]]} lines {text.splitlines()[int(bad[0])] if bad.size else b''}"
    return got


def test_txt_stream_semantics(pcp, oracle):
    """Missing fields keep the previous line's value, bad fields become 0 and end the line, empty lines repeat the
    point, overflow stores FLT_MAX, CRLF, tabs, signs, exponents, a last line without newline."""
    text = (b"1 2 3 4\n5 6 7\n\n8.5 abc 9 1\n 1e3 .5 -.25e-2 +7\r\n1e 2 3 4\n16777217 16777217.0000000000000001 1e39 5\n"
            b"\t-0.0 +.0e5 00012.500 1E+2\n- 1 2 3\n. 1 2 3\n1.5abc 2 3 4\n1..2 3 4 5\n1e5e3 4 5 6\n0x10 2 3 4\n1e+ 2 3 4\n"
            b"nan 1 2 3\ninf 1 2 3\n1,5 2 3 4\n   \n7 8 9 10 11 12\n-1e-50 1e-46 1.4e-45 7.1e-46\n3.4028236e38 3.4028235e38 -3.5e38 1\n3")
    got = _check(pcp, oracle, text)
    assert got.shape[0] == text.count(b"\n") + 1
    _check(pcp, oracle, b"1 2 3 4\n")          # trailing newline adds no line
    _check(pcp, oracle, b"\n\n\n")
    _check(pcp, oracle, b"5")
    _check(pcp, oracle, b"1 2 3 4\n\n")


def test_txt_rounding_hard_cases(pcp, oracle):
    """Halfway cases, denormals, 17-19 digit mantissas, big and small exponents: correctly rounded like strtof."""
    rs = np.random.default_rng(3)
    lines = []
    f = rs.integers(0, 0x7F7FFFFF, 20000, dtype=np.uint32).view(np.float32).astype(np.float64)
    for v in f[:5000]:
        lines.append("%.9g %.17g %.6e %.3f" % (v, v, v, v if abs(v) < 1e15 else 1.0))
    # exact midpoints between adjacent floats and their neighbours one digit away (printed exactly)
    for u in rs.integers(0x00800000, 0x7F000000, 3000, dtype=np.uint32):
        a = float(np.uint32(u).view(np.float32))
        b = float(np.uint32(u + 1).view(np.float32))
        mid = (a + b) / 2.0  # exact in double
        lines.append("%.18e %.17e %.12e %.9e" % (mid, mid, mid, mid))
    for u in rs.integers(1, 0x00800000, 2000, dtype=np.uint32):  # denormals
        a = float(np.uint32(u).view(np.float32))
        lines.append("%.18e %.10e %.8e %.15e" % (a, a * 1.0000001, a, a))
    for _ in range(3000):
        m = int(rs.integers(1, 10 ** 18))
        e = int(rs.integers(-60, 25))
        lines.append("%de%d %d.%de%d -%d 0.%018d" % (m, e, m % 1000, m % 10 ** 15, e // 2, m, m))
    text = ("\n".join(lines) + "\n").encode()
    _check(pcp, oracle, text)


def test_txt_long_mantissas(pcp, oracle):
    """More than 19 significant digits: the value rounds through its 19-digit prefix + a sticky bit — equal to strtof
    unless the prefix and its successor straddle a rounding boundary (then 1 ulp off; needs an adversarial input)."""
    rs = np.random.default_rng(4)
    v = rs.integers(0x00800000, 0x7F000000, 8000, dtype=np.uint32).view(np.float32).astype(np.float64) * (1.0 + rs.uniform(0, 1e-3, 8000))
    text = "".join("%.25e %.30e %.21e %.40e\n" % (a, a, a, a) for a in v).encode()
    ref = oracle.txt2pc(text)
    got = pcp.set_cloud_txt(text)
    np.testing.assert_array_equal(got.view(np.uint32), ref.view(np.uint32))
    # the adversarial case: an exact midpoint written out in full (> 19 digits) may land 1 ulp off, never more
    u = rs.integers(0x30000000, 0x50000000, 2000, dtype=np.uint32)
    mids = (u.view(np.float32).astype(np.float64) + (u + 1).view(np.float32).astype(np.float64)) / 2.0
    text = "".join("%.60e 1 2 3\n" % m for m in mids).encode()
    ref = oracle.txt2pc(text)
    got = pcp.set_cloud_txt(text)
    d = np.abs(got[:, 0].view(np.int32).astype(np.int64) - ref[:, 0].view(np.int32).astype(np.int64))
    assert d.max() <= 1


def test_txt_cloud_file_then_pca(pcp, oracle):
    """A cloud written like data/playground_subsampled.txt ("x y z intensity", src/smoothing.cpp:51-59) goes from
    text to k-NN PCA records without the points visiting the host."""
    from parity import check_features
    pts = synth.urban(23, 20000, extent=30.0)
    text = "".join("%.3f %.3f %.3f %.0f\n" % tuple(p) for p in pts.astype(np.float64)).encode()
    got = _check(pcp, oracle, text)
    pcp.build_grid(0.0, 20)
    rec = pcp.knn_pca(20)
    check_features(rec, oracle.knn_pca_cloud(got, 20, 0), None, "txt cloud k=20")


XYZ_CASES = [
    b"1 2 3 a\n4.5 5 6 zz\n7 8\n\n9 x 1 2\n",            # token's first byte; short lines keep old fields; matching failure ends a line
    b"1 2 3 7",                                            # no final newline: no repeated last point
    b"",                                                   # an empty file pushes one zero point
    b"\n",                                                 # ... and a lone newline two
    b"1e40 -1e40 1e-50 \xff\n",                            # +-inf, underflow to 0, a byte above 127 is a negative char
    b"  -0.5\t+.25 5. 0x\r\n1e 2 3 4\n1e+ 2 3 4\n. 1 2 3\n- 1 2 3\n+ 1 2 3\n1.5e3 2E-2 3e+1 q rest of line\n",
    b"1 2 3\n4 5 6 \n7 8 9 k\n1,2 3 4 5\n1 2 3abc 4\n00012.50 -0 1e0 9\n",
]


@pytest.mark.parametrize("text", XYZ_CASES)
def test_xyz_cases(pcp, oracle, text):
    """preprocess::xyz2pc (src/smoothing.cpp:83-125): the device parser against the reference's own getline + sscanf
    statements run on this image's C library (oracle/textio.cpp), bit for bit."""
    ref = oracle.xyz2pc(text)
    got = pcp.set_cloud_xyz(text)
    assert got.shape == ref.shape
    np.testing.assert_array_equal(got.view(np.uint32), ref.view(np.uint32))


def test_xyz_random_and_limits(pcp, oracle):
    rs = np.random.default_rng(11)
    pts = synth.urban(29, 30000, extent=40.0).astype(np.float64)
    lines = []
    for i, p in enumerate(pts):
        kind = rs.integers(0, 12)
        if kind == 0:
            lines.append("%.4f %.4f" % (p[0], p[1]))                      # short line
        elif kind == 1:
            lines.append("")                                              # empty line
        elif kind == 2:
            lines.append("%.9e %.3f %.17g %s" % (p[0], p[1], p[2], "ab"[i & 1]))
        elif kind == 3:
            lines.append("nope %.3f %.3f 1" % (p[1], p[2]))               # first field does not parse
        else:
            lines.append("%.3f %.3f %.3f %d" % (p[0], p[1], p[2], int(p[3])))
    text = ("\n".join(lines) + "\n").encode()
    ref = oracle.xyz2pc(text)
    got = pcp.set_cloud_xyz(text)
    np.testing.assert_array_equal(got.view(np.uint32), ref.view(np.uint32))
    assert got.shape[0] == len(lines) + 1  # the eof() loop's extra point
    # the parsed points are the current cloud
    pcp.build_grid(0.5, 0)
    assert pcp.radius_count(0.5, 4).shape[0] == got.shape[0]
    # a line the reference never returns from is an error here, as in the oracle
    # (getline(line, 256) stores up to 255 characters and tests the delimiter before the count: 255 fit, 256 set failbit)
    for tail in (b"\n", b""):
        bad = b"1 2 3 4\n" + b"7" * 256 + tail
        assert oracle.xyz2pc(bad) is None
        with pytest.raises(Exception):
            pcp.set_cloud_xyz(bad)
        for ok in (b"1 2 3 4\n" + b" " * 250 + b"7 8\n" + tail,          # 254 characters
                   b"1 2 3 4\n" + b"7" * 255 + tail,                      # 255 characters: the longest line the reference reads
                   b"1 2 3 4\n" + b" " * 248 + b"7.5 8 9" + tail):
            np.testing.assert_array_equal(pcp.set_cloud_xyz(ok).view(np.uint32), oracle.xyz2pc(ok).view(np.uint32))
