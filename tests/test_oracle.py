"""The CPU oracle against its pins: FIPS 180-4 known answers, hashlib, and itself."""
import hashlib

import numpy as np
import pytest

from oracle import sha256_oracle as oracle


def test_fips_known_answers(kat):
    for v in kat["fips"]:
        msg = v["unit"].encode() * v["repeat"]
        assert hashlib.sha256(msg).hexdigest() == v["sha256"]
        assert oracle.c_digest(msg).hex() == v["sha256"]


def test_padding_boundaries(kat):
    for v in kat["zeros"]:
        msg = bytes(v["len"])
        assert oracle.nblocks(v["len"]) == v["nblocks"]
        assert oracle.c_digest(msg).hex() == v["sha256"] == hashlib.sha256(msg).hexdigest()


def test_survey_appendix_values(kat):
    by_len = {v["len"]: v["sha256"] for v in kat["zeros"]}
    assert by_len[0] == "e3b0c44298fc1c149afbf4c8996fb92427ae41e4649b934ca495991b7852b855"
    assert by_len[55] == "02779466cdec163811d078815c633f21901413081449002f24aa3e80f0b88ef7"
    assert by_len[56] == "d4817aa5497628e7c77e6b606107042bbba3130888c5f47a375e6179be789fbb"
    assert by_len[64] == "f5a5fd42d16a20302798ef6ed309979b43003d2320d9f0e8ea9831a92759fb4b"
    assert by_len[4096] == "ad7facb2586fc6e966c004d7d1d16b024f5805ff7cb47c7a85dabd8b48892ca7"
    assert by_len[65536] == "de2f256064a0af797747c2b97505dc0b9f3df0de4f489eac731c23ae9ca9cc31"


def test_c_port_matches_hashlib_random_lengths():
    rng = np.random.default_rng(1234)
    lengths = np.concatenate([np.arange(0, 200), rng.integers(0, 1 << 17, size=200)])
    blob = rng.integers(0, 256, size=int(lengths.sum()), dtype=np.uint8)
    off = np.concatenate([[0], np.cumsum(lengths)[:-1]]).astype(np.uint64)
    got = oracle.c_digests_packed(blob, off, lengths.astype(np.uint64))
    got_mt = oracle.c_digests_packed(blob, off, lengths.astype(np.uint64), threads=4)
    for i, (o, n) in enumerate(zip(off, lengths)):
        want = hashlib.sha256(blob[int(o):int(o) + int(n)].tobytes()).digest()
        assert got[i].tobytes() == want
        assert got_mt[i].tobytes() == want


def test_synthetic_generator_golden(kat):
    for v in kat["synthetic"]:
        data = oracle.synth_object(v["seed"], v["object"], v["len"])
        assert len(data) == v["len"]
        assert data[:16].hex() == v["first16"]
        assert hashlib.sha256(data).hexdigest() == v["sha256"]


def test_synthetic_numpy_equals_c():
    for length, stride in ((1, 1), (7, 16), (100, 128), (4096, 4096), (1000, 1024)):
        buf = oracle.synth_batch(0xB200, 9, 5, length, stride)
        for i in range(5):
            assert bytes(buf[i * stride:i * stride + length]) == oracle.synth_object(0xB200, 9 + i, length)


def test_hashlib_mt_matches():
    objs = [oracle.synth_object(3, i, 3000 + i) for i in range(37)]
    assert oracle.hashlib_digests_mt(objs, 4) == oracle.hashlib_digests(objs)
