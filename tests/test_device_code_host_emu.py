"""Runs the product's device headers (fe.cuh / ec.cuh / hash.cuh / scan_core.cuh) compiled for the host with
VG_HOST_EMU (tests/host_emu/emu.cpp) against the oracle, so field arithmetic, the digests of all formats and the
batch-inversion walk (including windows that touch scalar 0 / n) are checked on a machine without a GPU.
This is a test harness only: the product never runs this code path on the CPU."""
import ctypes as C
import os
import random
import subprocess

import pytest

import oracle

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "host_emu", "emu.cpp")
LIB = os.path.join(HERE, "host_emu", "libvgen_emu.so")
P = 2**256 - 2**32 - 977
N = 0xFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFFEBAAEDCE6AF48A03BBFD25E8CD0364141
P_FIELD = 2**256 - 2**32 - 977


# The device headers are compiled for the host twice: as every format but P2TR builds them, and with the switches scan_fmt.inc sets
# for the P2TR translation unit (36-product squarings inside the inversion, the reduction's odd-limb products in their own pairs).
@pytest.fixture(scope="module", params=["default", "p2tr-build"])
def emu(request):
    csrc = os.path.join(HERE, "..", "vgen_b200", "csrc")
    deps = [SRC] + [os.path.join(csrc, f) for f in ("fe.cuh", "ec.cuh", "hash.cuh", "scan_core.cuh", "match.cuh", "taproot.cuh", "scan_types.h", "kernels.h")]
    deps = [d for d in deps if os.path.exists(d)]
    lib, flags = (LIB, []) if request.param == "default" else (LIB.replace(".so", "_p2tr.so"), ["-DVG_INV_FAST_SQR=1", "-DVG_REDUCE_SPLIT=1"])
    if not os.path.exists(lib) or any(os.path.getmtime(d) > os.path.getmtime(lib) for d in deps):
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-x", "c++", "-fPIC", "-shared"] + flags + ["-o", lib, SRC])
    return C.CDLL(lib)


def limbs(v):
    return (C.c_uint32 * 8)(*[(v >> (32 * i)) & 0xffffffff for i in range(8)])


def val(a):
    return sum(int(a[i]) << (32 * i) for i in range(8))


def test_field_ops(emu):
    rng = random.Random(42)
    edge = [0, 1, 2, P - 1, P, P + 1, 2**256 - 1, 2**256 - 2**32 - 978, 2**255, 977, 2**32 + 977, 2**256 - 2**33]
    vals = edge + [rng.getrandbits(256) for _ in range(300)]
    out = (C.c_uint32 * 8)()
    for i in range(0, len(vals) - 1):
        a, b = vals[i], vals[(i * 7 + 3) % len(vals)]
        emu.emu_fe_mul(limbs(a), limbs(b), out); assert val(out) == a * b % P
        emu.emu_fe_add(limbs(a), limbs(b), out); assert val(out) == (a + b) % P
        emu.emu_fe_sub(limbs(a), limbs(b), out); assert val(out) == (a - b) % P
        emu.emu_fe_sqr_fast(limbs(a), out); assert val(out) == a * a % P, hex(a)
    # squaring: carry patterns (all-ones limbs, single limbs, alternating)
    for a in [2**256 - 1, 2**224 - 1, (2**32 - 1) << 96, sum((2**32 - 1) << (64 * i) for i in range(4)), sum((2**32 - 1) << (64 * i + 32) for i in range(4)),
              0xFFFFFFFF00000001FFFFFFFF00000001FFFFFFFF00000001FFFFFFFF00000001] + [rng.getrandbits(256) | (2**256 - 2**200) for _ in range(50)]:
        emu.emu_fe_sqr_fast(limbs(a), out); assert val(out) == a * a % P, hex(a)
    for a in edge[1:4] + vals[12:40]:
        if a % P == 0:
            continue
        emu.emu_fe_inv(limbs(a), out)
        assert val(out) * a % P == 1


def test_scalar_mul_vs_oracle(emu):
    rng = random.Random(1)
    x = (C.c_uint32 * 8)(); y = (C.c_uint32 * 8)()
    for k in [1, 2, 3, N - 1, N - 2, 2**64, 2**255] + [rng.randrange(1, N) for _ in range(20)]:
        assert emu.emu_scalar_mul(limbs(k), x, y) == 0
        assert (val(x), val(y)) == oracle.pubkey(k)
    for k in [0, N, N + 1, 2**256 - 1]:
        assert emu.emu_scalar_mul(limbs(k), x, y) != 0


def test_windowed_scalar_mul_vs_oracle(emu):
    """ec.cuh scalar_mul_G (8-bit windows: centre set-up and degenerate-batch path on the device) == the oracle's k*G."""
    rng = random.Random(3)
    x = (C.c_uint32 * 8)(); y = (C.c_uint32 * 8)()
    ks = [1, 2, 255, 256, 257, 2**8 - 1, 2**64, 2**255, N - 1, N - 2, 0xFF00FF00FF00FF00FF00FF00FF00FF00FF00FF00FF00FF00FF00FF00FF00FF00 % N,
          (N + 1) // 2] + [rng.randrange(1, N) for _ in range(30)]
    for k in ks:
        assert emu.emu_scalar_mul_w8(limbs(k), x, y) == 0
        assert (val(x), val(y)) == oracle.pubkey(k), hex(k)
    for k in [0, N, 2**256 - 1]:
        assert emu.emu_scalar_mul_w8(limbs(k), x, y) != 0


@pytest.mark.parametrize("fmt", [0, 1, 2, 3, 5])
def test_digests_vs_golden(emu, golden, fmt):
    names = ["p2pkh", "p2pkh-uncompressed", "p2wpkh", "p2sh-p2wpkh", "p2tr", "ethereum"]
    dg = (C.c_uint32 * 8)()
    for v in golden["vectors"][:30]:
        emu.emu_digest(fmt, limbs(int(v["x"], 16)), limbs(int(v["y"], 16)), dg)
        got = b"".join(int(w).to_bytes(4, "little") for w in dg)[:20]
        assert got.hex() == v[names[fmt]]["digest"]


@pytest.mark.parametrize("fmt,start,T,H,nw", [
    (0, 1, 4, 8, 2),                 # starts at scalar 1
    (0, 0, 3, 4, 2),                 # touches scalar 0 -> degenerate batch flagged
    (0, 2**64, 5, 16, 3),
    (5, 2**64 + 12345, 2, 8, 2),
    (3, 2**200 + 7, 2, 4, 2),
    (1, 2**32 - 5, 2, 4, 2),
    (0, N - 40, 2, 8, 2),            # runs over n-1 -> degenerate
])
def test_walk_vs_oracle(emu, fmt, start, T, H, nw):
    total = T * 2 * H * nw
    out = C.create_string_buffer(20 * total)
    degen = emu.emu_walk_dump(fmt, limbs(start), T, H, nw, 20, out)
    want = oracle.digest_batch(fmt, start, total)
    B = 2 * H
    bad_batches = set()
    for i in range(total):
        k = start + i
        if k >= N:
            continue          # the engine clamps the range to n-1 (P.limit); the emulation harness has no limit
        g, w = out.raw[20 * i:20 * i + 20], want[32 * i:32 * i + 20]
        if g != w:
            bad_batches.add(i // B)
    # every mismatch must lie in a batch the walk flagged as degenerate (the engine redoes those with the slow kernel)
    assert len(bad_batches) <= degen
    if 0 < start and start + total + T * B * 2 < N:
        assert degen == 0 and not bad_batches
    else:
        assert degen > 0


def test_p2tr_two_stage_vs_oracle(emu):
    """taproot.cuh: stage 1 (walk -> even-Y internal keys) + stage 2 (TapTweak hash, t*G table addition, grouped
    inversion) must give the oracle's witness program (src/address.rs:136-140) for every key."""
    T, H, nw = 4, 8, 2                      # window = 64 keys: a multiple of VG_TAP_GROUP
    total = T * 2 * H * nw
    for start in (1, 2**64 + 5, N - 2 * total - 100):
        out = C.create_string_buffer(32 * total)
        degen = emu.emu_walk_dump(4, limbs(start), T, H, nw, 32, out)
        assert degen == 0
        want = oracle.digest_batch("p2tr", start, total)
        assert out.raw == want[:32 * total], start


def test_p2tr_table_geometry_words(emu):
    """scan_types.h: the geometry word of the P2TR fixed-base table — plain widths and signed mixed-radix windows — gives the
    window count, the table size and windows that cover exactly (signed) or at least (plain) the 256 scalar bits."""
    SIGNED = 0x40000000
    out = (C.c_ulonglong * 4)()
    for b in range(4, 27):
        emu.emu_tap_geom(b, out)
        n = (256 + b - 1) // b
        assert list(out) == [1, n, n << b, n * b]
    for W, H in [(17, 1), (25, 6), (28, 4), (4, 0), (5, 1), (7, 4), (3, 1), (21, 4)]:
        emu.emu_tap_geom(SIGNED | (H << 8) | W, out)
        n = (256 - H) // W
        assert list(out) == [1, n, H * (2**W + 1) + (n - H) * (2**(W - 1) + 1), 256], (W, H)
    emu.emu_tap_geom(SIGNED | (4 << 8) | 28, out)
    assert out[1] == 9 and out[2] == 1_744_830_473 and out[2] * 64 == 111_669_150_272                 # the 9-window table of the longest jobs: 111.7 GB
    for g in (0, 3, 27, 32, SIGNED | 28, SIGNED | (3 << 8) | 28, SIGNED | (1 << 8) | 2, SIGNED | (1 << 16) | (4 << 8) | 28, SIGNED | (20 << 8) | 31, SIGNED | (16 << 8) | 30):
        emu.emu_tap_geom(g, out)
        assert out[0] == 0, hex(g)


def test_p2tr_scalar_arithmetic_mod_n(emu):
    """taproot.cuh vg_scalar_add_mod_n / vg_scalar_neg_n (s = t + k or t + n - k mod n): the cases hashing never produces —
    sums that reach n exactly, pass 2^256, or come out 0 — and random operands, against Python integers."""
    rng = random.Random(341)
    pairs = [(0, 0), (N - 1, 1), (N - 1, N - 1), (1, N - 2), (N - 2**128, 2**128), (2**255, 2**255), (N - 1, 2**256 - N),
             (2**256 - N, 2**256 - N), (N // 2, N // 2), (N // 2, N // 2 + 1), (0, N - 1)]
    pairs += [(rng.randrange(N), rng.randrange(N)) for _ in range(2000)]
    pairs += [(N - 1 - rng.randrange(2**40), N - 1 - rng.randrange(2**40)) for _ in range(200)]
    for a, b in pairs:
        r = (C.c_uint32 * 8)()
        emu.emu_scalar_add_mod_n(limbs(a), limbs(b), r)
        assert val(r) == (a + b) % N, (hex(a), hex(b))
    for k in [1, 2, N - 1, N - 2, N // 2, 2**255, 2**128] + [rng.randrange(1, N) for _ in range(500)]:
        r = (C.c_uint32 * 8)()
        emu.emu_scalar_neg_n(limbs(k), r)
        assert val(r) == N - k


def test_p2tr_stage2_invalid_points_and_small_tweak_table(emu):
    """Stage 2 alone: 64 consecutive scalars from several bases, their public keys as stage 1 stores them (the scalar form reads
    x, the parity of y and derives the key's scalar from the window base; the older form wants the even-Y point). All-zero points
    (invalid scalar / outside the range) give all-zero programs; 8-, 5- and 7-bit windows agree with 4-bit ones."""
    scalar_form = emu.emu_tap_scalar_form()
    n = 64                                   # a multiple of VG_TAP_GROUP
    for base, dropped in [(1, {3, 9, 10}), (2**200 + 9, set()), (N - 40, {0}), (2**255 - 17, {63}), (7**80, {5, 6, 7, 8})]:
        ks = [base + i for i in range(n)]
        pts = (C.c_uint32 * (16 * n))()
        live = []
        for i, k in enumerate(ks):
            if 0 < k < N and i not in dropped:
                x, y = oracle.pubkey(k)
                if not scalar_form and (y & 1):
                    y = P_FIELD - y
                for l in range(8):
                    pts[16 * i + l] = (x >> (32 * l)) & 0xffffffff
                    pts[16 * i + 8 + l] = (y >> (32 * l)) & 0xffffffff
                live.append(i)
        assert live
        SIGNED = 0x40000000
        # 5 and 7: digits straddle 32-bit limbs and the top window is partial (as with 24-bit windows). Signed geometries
        # (scan_types.h: W | n_hi << 8; widths add up to 256): 64 x 4 bits, 1 x 6 + 50 x 5, 4 x 8 + 32 x 7 (the shape of the 9-window table)
        for bits in (4, 8, 5, 7, SIGNED | 4, SIGNED | (1 << 8) | 5, SIGNED | (4 << 8) | 7):
            out = C.create_string_buffer(32 * n)
            emu.emu_tap_stage2(pts, limbs(base), n, bits, out)
            for i, k in enumerate(ks):
                want = oracle.digest_batch("p2tr", k, 1)[:32] if i in live else b"\0" * 32
                assert out.raw[32 * i:32 * i + 32] == want, (bits, hex(k))


@pytest.mark.parametrize("fmt,pats", [
    ("p2pkh", ["abc$", "^1Cat", "[0-9]{3}$", "^1.*Z", ".", "^11", "^1[a-z]+[A-Z]", "(?i)dog", "^1111", "x{2,}"]),
    ("p2sh-p2wpkh", ["^3A", "aa$", "^3[0-9]", "(?i)cat"]),
    ("p2wpkh", ["dead$", "^bc1qq", "^bc1q.*zz", "aaa", "^bc1q[02-9]+", "q$"]),
    ("ethereum", ["^0xdead", "^0xDEAD", "(?i)^0xdead", "F00$", "^0x[0-9]{4}", "[A-F]{5}", "^0x00", "aB"]),
    ("p2pkh-uncompressed", ["^1A", "zz$"]),
    ("p2tr", ["^bc1pq", "dead$", "^bc1p.*zz", "aaa", "q$", "^bc1p[02-9]+"]),
])
def test_device_matcher_vs_host_regex(emu, fmt, pats):
    """match.cuh (on-device Base58Check / Bech32 / EIP-55 encoders feeding the DFA) must agree with the host's
    regex over the address string (Pattern::matches, src/pattern.rs:43-45) on every digest."""
    import vgen_b200 as vg
    rng = random.Random(5)
    F = vg.AddressFormat.parse(fmt)
    dl = 32 if fmt == "p2tr" else 20
    digests = [bytes(rng.randrange(256) for _ in range(dl)) for _ in range(400)]
    if dl == 32:
        digests = [d for d in digests] + [b"\0" * 32, b"\xff" * 32]
    else:
        digests += [b"\0" * 20, b"\0" * 19 + b"\1", b"\xff" * 20, b"\0\0\0" + bytes(rng.randrange(256) for _ in range(17)),
                b"\0" + bytes(rng.randrange(256) for _ in range(19)), b"\0" * 10 + b"\xff" * 10]
    emu.emu_confirm.argtypes = [C.c_int, C.c_void_p, C.c_char_p, C.c_char_p, C.c_char_p, C.c_uint32, C.c_uint32, C.c_uint32]
    for pat in pats:
        p = vg.Pattern(pat)
        d = p.device_dfa(F)
        n_hit = 0
        for dg in digests:
            words = (C.c_uint32 * 8)(*[int.from_bytes(dg[4 * i:4 * i + 4].ljust(4, b"\0"), "little") if 4 * i < dl else 0 for i in range(8)])
            got = emu.emu_confirm(int(F), words, d["table"], d["accept"], d["sym_class"], d["n_classes"], d["start"], d["flags"])
            addr = oracle.address_from_digest(fmt, dg)
            want = oracle.regex_match(pat, addr)
            assert bool(got) == want, (fmt, pat, addr)
            n_hit += want


@pytest.mark.parametrize("fmt", ["p2pkh", "p2sh-p2wpkh", "p2wpkh", "p2tr"])
def test_encoder_tables_agree_with_stepwise_paths(emu, fmt):
    """The shared-memory encoder tables of the walk kernels (Base58 radix conversion by column sums, Bech32 checksum by
    XOR of table rows, match.cuh) against the long-division / stepwise-polymod paths and the oracle's address strings on
    many digests: emu_confirm runs both device paths and returns -1 when they disagree. The patterns look at the last
    characters (checksum-dependent) and at a class over the whole string, so every digit reaches the DFA."""
    import vgen_b200 as vg
    rng = random.Random(11)
    F = vg.AddressFormat.parse(fmt)
    dl = 32 if fmt == "p2tr" else 20
    digests = [bytes(rng.randrange(256) for _ in range(dl)) for _ in range(3000)]
    digests += [b"\0" * dl, b"\xff" * dl, b"\0" * (dl - 1) + b"\1", b"\0" * 4 + b"\xff" * (dl - 4), b"\x80" + b"\0" * (dl - 1)]
    pats = ["[a-m]$", "[2-9]{2}", "^bc1[a-y0-9]+$"] if fmt in ("p2wpkh", "p2tr") else ["[a-m]$", "[1-9]{2}", "^[1-9A-Za-x]+$", "^1{2,}"]
    emu.emu_confirm.argtypes = [C.c_int, C.c_void_p, C.c_char_p, C.c_char_p, C.c_char_p, C.c_uint32, C.c_uint32, C.c_uint32]
    for pat in pats:
        d = vg.Pattern(pat).device_dfa(F)
        n_hit = 0
        for dg in digests:
            words = (C.c_uint32 * 8)(*[int.from_bytes(dg[4 * i:4 * i + 4].ljust(4, b"\0"), "little") if 4 * i < dl else 0 for i in range(8)])
            got = emu.emu_confirm(int(F), words, d["table"], d["accept"], d["sym_class"], d["n_classes"], d["start"], d["flags"])
            want = oracle.regex_match(pat, oracle.address_from_digest(fmt, dg))
            assert got == int(want), (fmt, pat, dg.hex(), got)
            n_hit += want
        assert 0 < n_hit < len(digests) or pat.startswith("^1{2"), (fmt, pat, n_hit)


@pytest.mark.parametrize("fmt", ["p2wpkh", "p2tr"])
def test_segwit_checksum_suffix_prefilter(emu, fmt):
    """Literal-suffix patterns on segwit formats (dead$): the device tests the checksum bits first and runs the DFA only on
    keys that pass (vg_dfa_dev::chk_mask / chk_value, lowering.cpp). Short suffixes so that random digests hit; the verdict
    must equal the oracle's regex over the address string, for matching and non-matching keys alike."""
    import vgen_b200 as vg
    CH = "qpzry9x8gf2tvdw0s3jn54khce6mua7l"
    rng = random.Random(23)
    F = vg.AddressFormat.parse(fmt)
    dl = 32 if fmt == "p2tr" else 20
    digests = [bytes(rng.randrange(256) for _ in range(dl)) for _ in range(3000)] + [b"\0" * dl, b"\xff" * dl]
    emu.emu_confirm_chk.argtypes = [C.c_int, C.c_void_p, C.c_char_p, C.c_char_p, C.c_char_p] + [C.c_uint32] * 5
    # one pattern is the real suffix of a digest's address, so a 6-character case matches at least once
    full = oracle.address_from_digest(fmt, digests[7])[-6:]
    for pat in ["q$", "e$", full[-2:] + "$", full[-3:] + "$", full + "$"]:
        lit = pat[:-1]
        mask, value = (1 << (5 * len(lit))) - 1, 0
        for ch in lit:
            value = (value << 5) | CH.index(ch)
        d = vg.Pattern(pat).device_dfa(F)
        n_hit = 0
        for dg in digests:
            words = (C.c_uint32 * 8)(*[int.from_bytes(dg[4 * i:4 * i + 4].ljust(4, b"\0"), "little") if 4 * i < dl else 0 for i in range(8)])
            got = emu.emu_confirm_chk(int(F), words, d["table"], d["accept"], d["sym_class"], d["n_classes"], d["start"], d["flags"], mask, value)
            want = oracle.regex_match(pat, oracle.address_from_digest(fmt, dg))
            assert got == int(want), (fmt, pat, dg.hex())
            n_hit += want
        assert n_hit > 0, (fmt, pat)


@pytest.mark.parametrize("fmt", ["p2pkh", "p2sh-p2wpkh"])
def test_base58_literal_suffix_residue_prefilter(emu, fmt):
    """Literal-suffix patterns on Base58 formats (Cat$): the last k characters are the Base58Check payload mod 58^k, so the device
    compares that residue (13 multiply-adds + one Barrett reduction, vg_dfa_dev::b58_mod / b58_vals) before the radix
    conversion and the DFA. The lowering exposes the residues (Pattern.lowering()["residue"]); the verdict must equal the
    oracle's regex over the address string, for matching and non-matching keys alike."""
    import vgen_b200 as vg
    B58 = "123456789ABCDEFGHJKLMNPQRSTUVWXYZabcdefghijkmnopqrstuvwxyz"
    rng = random.Random(31)
    F = vg.AddressFormat.parse(fmt)
    digests = [bytes(rng.randrange(256) for _ in range(20)) for _ in range(3000)]
    digests += [b"\0" * 20, b"\xff" * 20, b"\0" * 19 + b"\1", b"\0" * 10 + b"\xff" * 10]
    emu.emu_confirm_b58res.argtypes = [C.c_int, C.c_void_p, C.c_char_p, C.c_char_p, C.c_char_p] + [C.c_uint32] * 5 + [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32]
    full = oracle.address_from_digest(fmt, digests[5])
    for pat, ci in [("1$", False), ("z$", False), (full[-2:] + "$", False), (full[-3:] + "$", False), (full[-5:] + "$", False),
                    (full[-2:].swapcase() + "$", True), ("a$", True), ("l$", True), ("o$", True)]:
        low = vg.Pattern(pat, ci).lowering(F)
        assert low["kind"] == "dfa" and low["residue"], (pat, low)
        M, vals = low["residue"][0], list(low["residue"][1:])
        lit = pat[:-1]
        assert M == 58 ** len(lit)
        # every spelling the alphabet has, as a base-58 number
        import itertools
        spell = [sorted({c, c.swapcase()} & set(B58)) if ci else [c] for c in lit]
        want_vals = sorted(sum(B58.index(ch) * 58 ** (len(t) - 1 - i) for i, ch in enumerate(t)) for t in itertools.product(*spell))
        assert sorted(vals) == want_vals, (pat, vals, want_vals)
        w = (C.c_uint32 * 13)(*[pow(2, 16 * j, M) for j in range(13)])
        va = (C.c_uint32 * 8)(*(vals + [0] * (8 - len(vals))))
        mu = (1 << 64) // M
        d = vg.Pattern(pat, ci).device_dfa(F)
        n_hit = 0
        for dg in digests:
            words = (C.c_uint32 * 8)(*[int.from_bytes(dg[4 * i:4 * i + 4], "little") if i < 5 else 0 for i in range(8)])
            got = emu.emu_confirm_b58res(int(F), words, d["table"], d["accept"], d["sym_class"], d["n_classes"], d["start"], d["flags"],
                                         M, len(vals), va, w, mu & 0xffffffff, mu >> 32)
            want = oracle.regex_match(("(?i)" if ci else "") + pat, oracle.address_from_digest(fmt, dg))
            assert got == int(want), (fmt, pat, dg.hex())
            n_hit += want
        assert n_hit > 0, (fmt, pat)
    # patterns that are not a plain literal suffix get no residue test
    for pat in ["[ab]c$", "a.c$", "abcdef$", "^1a$", "a|b$"]:
        assert vg.Pattern(pat).lowering(F)["residue"] is None, pat


def test_eth_literal_suffix_prefilter(emu):
    """Ethereum literal hex suffixes (F00$, 12$): the device compares the last digest bits with the lower-cased literal before
    the DFAs and the EIP-55 hash. The verdict must equal the oracle's regex over the checksummed address."""
    import vgen_b200 as vg
    rng = random.Random(29)
    F = vg.AddressFormat.parse("ethereum")
    digests = [bytes(rng.randrange(256) for _ in range(20)) for _ in range(3000)]
    digests += [d[:18] + bytes.fromhex("af00") for d in digests[:200]] + [d[:16] + bytes.fromhex("deadbeef") for d in digests[:50]]
    emu.emu_confirm_chk.argtypes = [C.c_int, C.c_void_p, C.c_char_p, C.c_char_p, C.c_char_p] + [C.c_uint32] * 5
    for pat, ci in [("F00$", False), ("f00$", False), ("12$", False), ("a$", False), ("dEaDbEeF$", True), ("DEADBEEF$", False), ("00$", False)]:
        lit = pat[:-1]
        mask = (1 << (4 * len(lit))) - 1
        d = vg.Pattern(pat, ci).device_dfa(F)
        n_hit = 0
        for dg in digests:
            words = (C.c_uint32 * 8)(*[int.from_bytes(dg[4 * i:4 * i + 4], "little") if i < 5 else 0 for i in range(8)])
            got = emu.emu_confirm_chk(int(F), words, d["table"], d["accept"], d["sym_class"], d["n_classes"], d["start"], d["flags"], mask, int(lit, 16))
            want = oracle.regex_match(("(?i)" if ci else "") + pat, oracle.address_from_digest("ethereum", dg))
            assert got == int(want), (pat, dg.hex())
            n_hit += want
        assert n_hit > 0 or pat == "DEADBEEF$", (pat, n_hit)


def test_eth_folded_dfa_prefilter(emu):
    """match.cuh vg_match_eth with the case-folded DFA appended (lowering.cpp): the prefilter on the lower-case hex must
    never reject an address the case-sensitive pattern matches, and the combined verdict must equal the host regex."""
    import vgen_b200 as vg
    rng = random.Random(9)
    F = vg.AddressFormat.Ethereum
    emu.emu_confirm_eth_fold.restype = C.c_int
    digests = [bytes(rng.randrange(256) for _ in range(20)) for _ in range(600)] + [b"\0" * 20, b"\xff" * 20, b"\xab" * 20]
    digests += [bytes(rng.randrange(256) for _ in range(18)) + bytes([rng.randrange(16) * 16 + 15, 0]) for _ in range(80)]   # ...f00
    for pat in ["F00$", "dEaD", "[A-F]{3}$", "aB", "^0xDe", "(?-i:Ab)", "^0x[a-f]", "c0FFee|Bad", "e$"]:
        d, f = vg.Pattern(pat).device_dfa(F), vg.Pattern(pat.replace("(?-i:", "(?:"), True).device_dfa(F)
        assert d["flags"] & 1 and not f["flags"] & 1
        table, accept = d["table"] + f["table"], d["accept"] + f["accept"]
        n_hit = 0
        for dg in digests:
            words = (C.c_uint32 * 8)(*[int.from_bytes(dg[4 * i:4 * i + 4], "little") if i < 5 else 0 for i in range(8)])
            got = emu.emu_confirm_eth_fold(words, table, accept, d["sym_class"], d["n_classes"], d["start"], d["flags"],
                                           len(d["table"]) // 2, len(d["accept"]), f["n_classes"], f["start"], f["sym_class"][:16])
            want = oracle.regex_match(pat, oracle.address_from_digest("ethereum", dg))
            assert bool(got) == want, (pat, dg.hex())
            n_hit += want
        assert pat not in ("F00$", "e$", "^0x[a-f]", "aB") or n_hit > 0, pat
