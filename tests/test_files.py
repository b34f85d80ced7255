"""CPU: the two binary file formats of the boundary (mkfhe_b200/csrc/keyfile.h) -- key file and ciphertext / wire-table file.
Writing and verifying need no GPU.  The reader must survive hostile files: truncation, wrong counts, duplicate or
unknown sections, kskA without kskB, a flipped byte -- a Python exception, never a crash.  The C++ bridge
(BinFHEBatchB200::SaveKeyFile over the reference's UniEncBTKey) writes the bytes the flat exports hold."""
import struct

import numpy as np
import pytest

import mkfhe_b200 as mk
from mkfhe_b200 import lib as mklib
from oracle import refapi
from util import load_golden


def _params(p):
    return mk.Params(k=p.k, N=p.N, Q=p.Q, psi=p.psi, q=p.q, qKS=p.qKS, n=p.n, baseG_log2=p.baseG_log2, digitsG=p.digitsG,
                     baseKS=p.baseKS, dKS=p.dKS, variant=p.variant, scheme=p.scheme, device=-1)


def parse(path):
    """The documented layout, restated: header (72 bytes) then sections of 24-byte descriptors + payload + pad to 16."""
    raw = open(path, "rb").read()
    magic, version, nsec = raw[:8], *struct.unpack_from("<II", raw, 8)
    par = struct.unpack_from("<13I", raw, 16)
    at, out = 72, {}
    for _ in range(nsec):
        tag, zero, words, fnv = struct.unpack_from("<IIQQ", raw, at)
        at += 24
        out[tag] = np.frombuffer(raw, dtype="<u4", count=words, offset=at).copy()
        at += 4 * words + (16 - (24 + 4 * words) % 16) % 16
    return magic, version, par, out


def test_keyfile_layout_is_as_documented(tmp_path):
    g = load_golden("toy_mklwe_k2_N128")
    P = _params(g["p"])
    path = str(tmp_path / "k.mkfhe")
    mk.SaveKeyFile(path, P, g["bk"], g["pkey"], kskA=g["kskA"], kskB=g["kskB"])
    magic, version, par, sec = parse(path)
    assert magic == b"MKFHEKEY" and version == 1
    assert par == (P.k, P.N, P.Q, P.psi, P.q, P.qKS, P.n, P.baseG_log2, P.digitsG, P.baseKS, P.dKS, P.variant, P.scheme)
    assert sorted(sec) == [1, 2, 4, 5]
    assert (sec[1] == g["bk"].ravel()).all() and (sec[2] == g["pkey"].ravel()).all()
    assert (sec[4] == g["kskA"].ravel()).all() and (sec[5] == g["kskB"].ravel()).all()


def _section_offsets(raw):
    nsec = struct.unpack_from("<I", raw, 12)[0]
    at, offs = 72, []
    for _ in range(nsec):
        tag, zero, words, fnv = struct.unpack_from("<IIQQ", raw, at)
        offs.append((at, tag, words))
        at += 24 + 4 * words + (16 - (24 + 4 * words) % 16) % 16
    return offs


@pytest.mark.parametrize("attack", ["truncate_payload", "truncate_header", "huge_words", "huge_count", "zero_words", "dup_tag", "bad_tag",
                                    "kska_without_kskb", "flip_byte", "flip_checksum", "wrong_magic", "empty"])
def test_hostile_key_files_are_refused(attack, tmp_path):
    g = load_golden("toy_mklwe_k2_N128")
    path = str(tmp_path / "k.mkfhe")
    mk.SaveKeyFile(path, _params(g["p"]), g["bk"], g["pkey"], kskA=g["kskA"], kskB=g["kskB"])
    mk.KeyFileInfo(path)  # sound before the attack
    raw = bytearray(open(path, "rb").read())
    offs = _section_offsets(raw)
    if attack == "truncate_payload":
        raw = raw[: offs[1][0] + 24 + 100]
    elif attack == "truncate_header":
        raw = raw[:40]
    elif attack == "huge_words":      # a section that claims 2^61 words: must not be allocated
        struct.pack_into("<Q", raw, offs[0][0] + 8, 1 << 61)
    elif attack == "huge_count":
        struct.pack_into("<I", raw, 12, 0xFFFFFFFF)
    elif attack == "zero_words":
        struct.pack_into("<Q", raw, offs[0][0] + 8, 0)
    elif attack == "dup_tag":
        struct.pack_into("<I", raw, offs[1][0], offs[0][1])
    elif attack == "bad_tag":
        struct.pack_into("<I", raw, offs[1][0], 77)
    elif attack == "kska_without_kskb":  # drop the last section (kskB) and fix the count
        raw = raw[: offs[3][0]]
        struct.pack_into("<I", raw, 12, 3)
    elif attack == "flip_byte":
        raw[offs[2][0] + 24 + 17] ^= 0x40
    elif attack == "flip_checksum":
        raw[offs[0][0] + 16] ^= 1
    elif attack == "wrong_magic":
        raw[:8] = b"MKFHECTX"
    elif attack == "empty":
        raw = bytearray()
    open(path, "wb").write(raw)
    with pytest.raises(mk.MkfheError):
        mk.KeyFileInfo(path)


def test_ciphertext_file_round_trip_and_refusals(tmp_path):
    rng = np.random.default_rng(12)
    for name in ("toy_mklwe_k2_N128", "toy_mkntru_k4_N256"):
        g = load_golden(name)
        P = _params(g["p"])
        lwe = P.scheme == 1
        n = 37
        ct = rng.integers(0, P.q, size=(n, P.k, P.n), dtype=np.uint32)
        b = rng.integers(0, P.q, size=n, dtype=np.uint32) if lwe else None
        path = str(tmp_path / f"{name}.ct")
        mk.SaveCiphertexts(path, P, ct, b)
        Pr, ct2, b2 = mk.LoadCiphertexts(path, P)
        assert (Pr.k, Pr.n, Pr.q, Pr.scheme) == (P.k, P.n, P.q, P.scheme)
        assert (ct2 == ct).all() and ((b2 == b).all() if lwe else b2 is None)
        magic, version, par, sec = parse(path)
        assert magic == b"MKFHECTX" and sorted(sec) == ([8, 9] if lwe else [8])
        assert (sec[8] == ct.ravel()).all()
        other = _params(load_golden("toy_mklwe_k3_N32_B5")["p"])
        with pytest.raises(mk.MkfheError):   # parameters of another set
            mk.LoadCiphertexts(path, other)
        with pytest.raises(mk.MkfheError):   # a ciphertext file is not a key file
            mk.KeyFileInfo(path)
        raw = bytearray(open(path, "rb").read())
        raw[72 + 24 + 5] ^= 2
        open(path, "wb").write(raw)
        with pytest.raises(mk.MkfheError):   # checksum
            mk.LoadCiphertexts(path)
    g = load_golden("toy_mklwe_k2_N128")
    with pytest.raises(mk.MkfheError):       # MK-LWE ciphertexts carry b
        mk.SaveCiphertexts(str(tmp_path / "x.ct"), _params(g["p"]), g["g_ct1"], None)


@pytest.mark.skipif(not refapi.available(), reason="oracle/_ref/libmkfhe_ref.so not present")
@pytest.mark.parametrize("toy", [(0, 2, 27, 32, 4, 45181, 45181, 0.75, 32, 1 << 9, 32), (1, 2, 27, 32, 4, 32749, 32749, 1.9, 32, 1 << 9, 2)],
                         ids=["mkntru", "mklwe"])
def test_cpp_bridge_writes_the_reference_objects_as_a_key_file(toy, tmp_path):
    """integration/binfhe_batch_b200.h SaveKeyFile(params, method, UniEncBTKey, ctNAND), compiled against the unmodified
    reference: the file's sections are the reference's keys as the flat exports hold them.  Host only (no GPU)."""
    r = refapi.Ref(*toy)
    r.keygen()
    path = str(tmp_path / "ref.mkfhe")
    r.batch_b200_save_keyfile(mklib.LIB_PATH, path)
    P, sizes = mk.KeyFileInfo(path)
    assert (P.k, P.N, P.n, P.q, P.qKS, P.Q, P.psi, P.scheme) == (r.p.k, r.p.N, r.p.n, r.p.q, r.p.qKS, r.p.Q, r.p.psi, r.p.scheme)
    _, _, _, sec = parse(path)
    assert (sec[1] == r.bk()).all() and (sec[2] == r.pkey()).all()
    if r.p.scheme == 0:
        assert (sec[3] == r.ksk_mntru()[0]).all() and (sec[6] == r.ctnand()).all() and sorted(sec) == [1, 2, 3, 6]
    else:
        A, B = r.ksk_mklwe()
        assert (sec[4] == A).all() and (sec[5] == B).all() and sorted(sec) == [1, 2, 4, 5]
    r.close()
