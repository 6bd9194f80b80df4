"""Parity of the CUDA path with the oracle, through the C ABI, on a B200.

Bit-exact is the bar (integer/byte work): every digest is compared with hashlib.sha256
(the oracle BASELINE.json names) on the same bytes.  Every load path (per-lane loads,
per-lane TMA bulk copies, TMA 2-D tiles) and every FMA-pipe level is exercised.
"""
import hashlib

import numpy as np
import pytest
import torch

from bee_code_interpreter_fs_b200 import _ffi
from oracle import sha256_oracle as oracle

pytestmark = pytest.mark.gpu

KERNELS = [0, 1, 2]
FMA_LEVELS = [0, 1, 2, 3, 4]


def configure(h, kernel=None, fma=None, ctas=None):
    if kernel is not None:
        h.set_option(_ffi.OPT_KERNEL, kernel)
    if fma is not None:
        h.set_option(_ffi.OPT_FMA_LEVEL, fma)
    if ctas is not None:
        h.set_option(_ffi.OPT_CTAS_PER_SM, ctas)


@pytest.fixture(autouse=True)
def restore_defaults(hasher):
    saved = {k: hasher.get_option(k) for k in (_ffi.OPT_KERNEL, _ffi.OPT_FMA_LEVEL, _ffi.OPT_CTAS_PER_SM,
                                              _ffi.OPT_WAVE_BYTES, _ffi.OPT_LONG_MAX_OBJECTS, _ffi.OPT_LONG_MIN_BYTES)}
    yield
    for k, v in saved.items():
        hasher.set_option(k, v)


def hexes(digests):
    return [bytes(d).hex() for d in np.asarray(digests)]


# ------------------------------------------------------------ known answers --
@pytest.mark.parametrize("kernel", KERNELS)
@pytest.mark.parametrize("fma", FMA_LEVELS)
def test_fips_and_boundary_vectors_host_path(hasher, kat, kernel, fma):
    configure(hasher, kernel, fma)
    msgs = [v["unit"].encode() * v["repeat"] for v in kat["fips"]] + [bytes(v["len"]) for v in kat["zeros"]]
    want = [v["sha256"] for v in kat["fips"]] + [v["sha256"] for v in kat["zeros"]]
    assert hasher.hash_many(msgs) == want


@pytest.mark.parametrize("ctas", [2, 3, 4])
def test_register_budget_variants(hasher, kat, ctas):
    configure(hasher, ctas=ctas)
    for kernel in KERNELS:
        configure(hasher, kernel)
        msgs = [bytes(v["len"]) for v in kat["zeros"]]
        assert hasher.hash_many(msgs) == [v["sha256"] for v in kat["zeros"]]


# ------------------------------------------------- device path, uniform layout --
@pytest.mark.parametrize("kernel", KERNELS)
@pytest.mark.parametrize("length,stride,n", [
    (64, 64, 33), (4096, 4096, 1000), (65536, 65536, 70), (1000, 1008, 257), (55, 64, 64), (56, 64, 31),
    (119, 128, 5), (4097, 4112, 300), (1, 16, 100), (0, 16, 3), (129, 129, 40), (8192, 16384, 513),
])
def test_uniform_device_path_matches_oracle(hasher, kernel, length, stride, n):
    """Also proves the device generator emits the oracle's synthetic bytes."""
    configure(hasher, kernel)
    buf = torch.zeros(max(16, n * stride + 16), dtype=torch.uint8, device="cuda:0")
    hasher.gen_synthetic(buf, 0xB200, 11, n, length, stride)
    dig = hasher.digest_device_uniform(buf, stride, length, n)
    torch.cuda.synchronize()
    got = hexes(dig.cpu().numpy())
    want = [hashlib.sha256(oracle.synth_object(0xB200, 11 + i, length)).hexdigest() for i in range(n)]
    assert got == want
    if length:
        host = buf.cpu().numpy()
        assert bytes(host[stride * (n - 1): stride * (n - 1) + length]) == oracle.synth_object(0xB200, 11 + n - 1, length)


@pytest.mark.parametrize("n", [148 * 24 * 32 + 1, 120000, 148 * 28 * 32])
def test_one_wave_launch_shape_for_batches_just_over_a_wave(hasher, n):
    """Uniform batches of 1.0-1.17 waves (24 to 28 warps per SM) run as seven 128-thread CTAs per SM
    instead of three 256-thread ones; the window's edges and its middle, every digest."""
    assert hasher.device_info()["sm_count"] == 148
    length, stride = 100, 112
    buf = torch.empty(n * stride, dtype=torch.uint8, device="cuda:0")
    hasher.gen_synthetic(buf, 3, 0, n, length, stride)
    got = hasher.digest_device_uniform(buf, stride, length, n).cpu().numpy()
    host = buf.cpu().numpy().reshape(n, stride)[:, :length]
    want = np.frombuffer(b"".join(hashlib.sha256(row).digest() for row in host), dtype=np.uint8).reshape(n, 32)
    assert (got == want).all()
    configure(hasher, ctas=3)  # the ordinary shape on the same batch
    assert (hasher.digest_device_uniform(buf, stride, length, n).cpu().numpy() == want).all()


@pytest.mark.parametrize("fma", FMA_LEVELS)
def test_fma_levels_agree_on_tma_path(hasher, fma):
    configure(hasher, 2, fma)
    n, length = 2000, 4096
    buf = torch.empty(n * length, dtype=torch.uint8, device="cuda:0")
    hasher.gen_synthetic(buf, 5, 0, n, length)
    got = hexes(hasher.digest_device_uniform(buf, length, length, n).cpu().numpy())
    assert got == [hashlib.sha256(oracle.synth_object(5, i, length)).hexdigest() for i in range(n)]


# ------------------------------------------- device path, ragged offsets/lengths --
@pytest.mark.parametrize("kernel", KERNELS)
@pytest.mark.parametrize("align", [16, 1])
def test_ragged_device_path_any_alignment(hasher, kernel, align):
    configure(hasher, kernel)
    rng = np.random.default_rng(42 + align)
    lengths = np.concatenate([np.arange(0, 140), rng.integers(0, 1 << 14, size=400), [1 << 17, (1 << 17) + 3]])
    rng.shuffle(lengths)
    slots = (lengths + align - 1) // align * align + (0 if align > 1 else rng.integers(0, 3, size=lengths.size))
    off = np.concatenate([[0], np.cumsum(slots)[:-1]]).astype(np.uint64)
    if align == 1:
        off += 1  # nothing is aligned
    blob = rng.integers(0, 256, size=int(off[-1] + lengths[-1] + 64), dtype=np.uint8)
    d_blob = torch.from_numpy(blob).cuda()
    d_off = torch.from_numpy(off.astype(np.int64)).cuda()
    d_len = torch.from_numpy(lengths.astype(np.int64)).cuda()
    got = hexes(hasher.digest_device(d_blob, d_off, d_len).cpu().numpy())
    want = [hashlib.sha256(blob[int(o):int(o) + int(n)].tobytes()).hexdigest() for o, n in zip(off, lengths)]
    assert got == want
    # longest-first schedule: same digests, object-major output
    order = torch.from_numpy(np.argsort(-lengths, kind="stable").astype(np.uint32).view(np.int32)).cuda()
    got2 = hexes(hasher.digest_device(d_blob, d_off, d_len, order=order).cpu().numpy())
    assert got2 == want


@pytest.mark.parametrize("per_warp", [1, 2, 4, 8, 16, 32])
@pytest.mark.parametrize("align", [16, 1])
def test_warp_per_object_kernel_any_length_any_alignment(hasher, align, per_warp):
    """cas_hash_device_long: schedules expanded by the warp's lanes, rounds serial, 1 to 32 objects
    per warp (the engine picks that from the object count; here it is forced through the
    warp budget); same digests for every mix of lengths in a warp."""
    hasher.set_option(_ffi.OPT_LONG_MAX_OBJECTS, max(1, 61 // per_warp))  # 61 objects below
    rng = np.random.default_rng(99 + align)
    lengths = np.array([0, 1, 55, 56, 63, 64, 65, 119, 120, 127, 128, 2047, 2048, 2049, 2111, 2112, 4095, 4096,
                        65536, 100001, (1 << 20) + 3] + rng.integers(0, 9000, size=40).tolist())
    slots = (lengths + align - 1) // align * align + (0 if align > 1 else rng.integers(0, 3, size=lengths.size))
    off = np.concatenate([[0], np.cumsum(slots)[:-1]]).astype(np.uint64)
    if align == 1:
        off += 3
    blob = rng.integers(0, 256, size=int(off[-1] + lengths[-1] + 64), dtype=np.uint8)
    d_blob = torch.from_numpy(blob).cuda()
    d_off = torch.from_numpy(off.astype(np.int64)).cuda()
    d_len = torch.from_numpy(lengths.astype(np.int64)).cuda()
    want = [hashlib.sha256(blob[int(o):int(o) + int(n)].tobytes()).hexdigest() for o, n in zip(off, lengths)]
    got = hexes(hasher.digest_device(d_blob, d_off, d_len, warp_per_object=True).cpu().numpy())
    assert got == want
    order = torch.from_numpy(np.argsort(-lengths, kind="stable").astype(np.int32)).cuda()
    got = hexes(hasher.digest_device(d_blob, d_off, d_len, order=order, warp_per_object=True).cpu().numpy())
    assert got == want


def test_few_long_uniform_objects_take_the_warp_kernel_and_agree_with_lanes(hasher):
    n, length = 40, (3 << 20) + 17
    stride = (length + 15) // 16 * 16
    buf = torch.empty(n * stride, dtype=torch.uint8, device="cuda:0")
    hasher.gen_synthetic(buf, 123, 0, n, length, stride)
    a = hexes(hasher.digest_device_uniform(buf, stride, length, n).cpu().numpy())  # warp per object (n small)
    saved = hasher.get_option(_ffi.OPT_LONG_MAX_OBJECTS)
    assert saved == 4 * hasher.device_info()["sm_count"]
    hasher.set_option(_ffi.OPT_LONG_MAX_OBJECTS, 0)
    try:
        b = hexes(hasher.digest_device_uniform(buf, stride, length, n).cpu().numpy())  # lane per object
    finally:
        hasher.set_option(_ffi.OPT_LONG_MAX_OBJECTS, saved)
    assert a == b
    assert a[0] == hashlib.sha256(oracle.synth_object(123, 0, length)).hexdigest()
    assert a[-1] == hashlib.sha256(oracle.synth_object(123, n - 1, length)).hexdigest()


# ------------------------------------------------------------ host path (e2e) --
@pytest.mark.parametrize("kernel", KERNELS)
def test_host_path_random_lengths_property(hasher, kernel):
    configure(hasher, kernel)
    hasher.set_option(_ffi.OPT_WAVE_BYTES, 1 << 20)  # force many waves and buffer reuse
    rng = np.random.default_rng(7)
    lengths = np.concatenate([rng.integers(0, 1 << 17, size=300), [0, 0, 1, 3 << 20]])
    objs = [rng.integers(0, 256, size=int(n), dtype=np.uint8).tobytes() for n in lengths]
    assert hasher.hash_many(objs) == [hashlib.sha256(o).hexdigest() for o in objs]


@pytest.mark.parametrize("pinned", [False, True])
def test_host_path_many_epochs_and_objects_larger_than_a_chunk(hasher, pinned):
    """Arena halves are reused across epochs and an object may span several copy chunks."""
    saved = hasher.get_option(_ffi.OPT_ARENA_BYTES)
    hasher.set_option(_ffi.OPT_ARENA_BYTES, 64 << 20)  # 32 MiB halves
    hasher.set_option(_ffi.OPT_WAVE_BYTES, 4 << 20)
    try:
        rng = np.random.default_rng(11)
        lengths = np.concatenate([[20 << 20, (9 << 20) + 7, 5 << 20], rng.integers(0, 300000, size=700)]).astype(np.uint64)
        rng.shuffle(lengths)
        off = np.concatenate([[0], np.cumsum(lengths)[:-1]]).astype(np.uint64)
        total = int(lengths.sum())
        assert total > 3 * (32 << 20)
        data = rng.integers(0, 256, size=total, dtype=np.uint8)
        if pinned:
            host = torch.empty(total, dtype=torch.uint8).pin_memory()
            host.numpy()[:] = data
            got = hasher.digest_packed(host, off, lengths)
        else:
            got = hasher.digest_packed(data, off, lengths)
        want = [hashlib.sha256(data[int(o):int(o) + int(n)].tobytes()).hexdigest() for o, n in zip(off, lengths)]
        assert hexes(got) == want
        # uniform batch over several epochs, straight from pinned memory
        n, length = 3000, 65536
        host = torch.empty(n * length, dtype=torch.uint8).pin_memory()
        host.numpy()[:] = oracle.synth_batch(21, 0, n, length)
        dig = hasher.digest_packed(host, np.arange(n, dtype=np.uint64) * length, np.full(n, length, dtype=np.uint64))
        assert hexes(dig) == [hashlib.sha256(oracle.synth_object(21, i, length)).hexdigest() for i in range(n)]
    finally:
        hasher.set_option(_ffi.OPT_ARENA_BYTES, saved)


def test_host_path_pinned_contiguous_and_mixed_types(hasher):
    n, length = 3000, 4096
    host = torch.empty(n * length, dtype=torch.uint8).pin_memory()
    host.numpy()[:] = oracle.synth_batch(9, 0, n, length)
    off = np.arange(n, dtype=np.uint64) * length
    dig = hasher.digest_packed(host, off, np.full(n, length, dtype=np.uint64))
    assert hexes(dig) == [hashlib.sha256(oracle.synth_object(9, i, length)).hexdigest() for i in range(n)]
    mixed = [b"abc", bytearray(b"defg"), memoryview(b"hij"), np.arange(10, dtype=np.uint8), b""]
    want = [hashlib.sha256(bytes(m)).hexdigest() for m in mixed]
    assert hasher.hash_many(mixed) == want
    assert hasher.hash_many([]) == []


def test_million_a_and_half_gib_message_use_64bit_length(hasher, kat):
    """The FIPS one-million-'a' vector, and a message just over 2^29 bytes (its bit length needs
    33 bits) through the warp-per-object kernel.  tests/test_gpu_configs.py sends a 2^30-byte
    object through the lane kernels as well."""
    assert hasher.hash_one(b"a" * 1_000_000) == kat["fips"][4]["sha256"]
    n = (1 << 29) + 5
    buf = torch.empty(n + 11, dtype=torch.uint8, device="cuda:0")
    hasher.gen_synthetic(buf, 77, 3, 1, n, n)
    got = hexes(hasher.digest_device_uniform(buf, n + 11, n, 1).cpu().numpy())[0]
    assert got == hashlib.sha256(buf[:n].cpu().numpy().tobytes()).hexdigest()


# ------------------------------------------ full-size, size-independent properties --
def test_full_size_4k_config_sampled_and_cross_checked(hasher):
    """BASELINE config 2 at full size (2^20 x 4 KiB = 4 GiB): a seeded sample of objects
    against the oracle, and a checksum of all digests equal across the three load paths."""
    n, length = 1 << 20, 4096
    buf = torch.empty(n * length, dtype=torch.uint8, device="cuda:0")
    hasher.gen_synthetic(buf, 0xB200, 0, n, length)
    sums = []
    for kernel in KERNELS:
        configure(hasher, kernel)
        dig = hasher.digest_device_uniform(buf, length, length, n)
        torch.cuda.synchronize()
        sums.append(hashlib.sha256(dig.cpu().numpy().tobytes()).hexdigest())
        if kernel == 2:
            host = dig.cpu().numpy()
            rng = np.random.default_rng(1)
            for i in [0, 1, 31, 32, 255, 256, n - 1] + rng.integers(0, n, size=500).tolist():
                assert bytes(host[i]).hex() == hashlib.sha256(oracle.synth_object(0xB200, i, length)).hexdigest()
    assert sums[0] == sums[1] == sums[2]


def test_full_size_64k_config_cross_checked(hasher):
    """BASELINE config 3 at full size (2^20 x 64 KiB = 64 GiB): the TMA-tile kernel and the
    per-lane-load kernel must agree on every one of the 2^20 digests (checksum of checksums),
    and a seeded sample is compared with the oracle."""
    free, _ = torch.cuda.mem_get_info(0)
    n, length = 1 << 20, 65536
    if free < n * length + (4 << 30):
        pytest.skip("not enough free device memory for the 64 GiB batch")
    buf = torch.empty(n * length, dtype=torch.uint8, device="cuda:0")
    hasher.gen_synthetic(buf, 0xB200, 0, n, length)
    sums = {}
    for kernel in (2, 0):
        configure(hasher, kernel)
        dig = hasher.digest_device_uniform(buf, length, length, n)
        torch.cuda.synchronize()
        host = dig.cpu().numpy()
        sums[kernel] = hashlib.sha256(host.tobytes()).hexdigest()
        if kernel == 2:
            rng = np.random.default_rng(2)
            for i in [0, 31, 32, 255, 256, n - 1] + rng.integers(0, n, size=250).tolist():
                assert bytes(host[i]).hex() == hashlib.sha256(oracle.synth_object(0xB200, i, length)).hexdigest()
    assert sums[2] == sums[0]
    del buf


def test_c_abi_rejects_bad_arguments_with_a_message(hasher):
    """Errors are codes plus text, never a placeholder digest (include/b200cas.h conventions)."""
    from bee_code_interpreter_fs_b200 import CasError

    buf = torch.zeros(4096, dtype=torch.uint8, device="cuda:0")
    with pytest.raises(CasError, match="stride smaller than length"):
        hasher.digest_device_uniform(buf, 16, 64, 4)
    with pytest.raises(CasError, match="CAS_OPT_FMA_LEVEL"):
        hasher.set_option(_ffi.OPT_FMA_LEVEL, 9)
    with pytest.raises(CasError, match="unknown option"):
        hasher.set_option(99, 1)
    with pytest.raises(ValueError):
        hasher.digest_packed(None, np.zeros(3, np.uint64), np.zeros(2, np.uint64))
    with pytest.raises(ValueError, match="too small"):
        hasher.gen_synthetic(buf, 1, 0, 100, 4096)
    lib = _ffi.load()
    assert lib.cas_hash_host(hasher._ctx, None, None, None, 5, None) == _ffi.CAS_E_ARG
    assert b"NULL" in lib.cas_last_error(hasher._ctx)
    assert lib.cas_hash_host(hasher._ctx, None, None, None, 0, None) == _ffi.CAS_OK  # empty batch is fine
    assert hasher.hash_many([b"still works"]) == [hashlib.sha256(b"still works").hexdigest()]


def test_digests_independent_of_batch_position_and_neighbours(hasher):
    """Idempotence / independence: an object's id does not depend on which lane or warp hashed it."""
    rng = np.random.default_rng(3)
    objs = [rng.integers(0, 256, size=int(n), dtype=np.uint8).tobytes() for n in rng.integers(0, 20000, size=257)]
    a = hasher.hash_many(objs)
    perm = rng.permutation(len(objs))
    b = hasher.hash_many([objs[i] for i in perm])
    assert [a[i] for i in perm] == b
    assert hasher.hash_many(objs) == a
