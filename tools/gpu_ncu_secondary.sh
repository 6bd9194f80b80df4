#!/bin/bash
# ncu full captures of the two secondary kernels (per-lane loads on a ragged batch, warp per
# object on a few long objects).  Run under gpurun, one GPU.  $1 = tag.
TAG=${1:-r01}
mkdir -p gpurun_out
cat > /tmp/secondary.py <<'PY'
import sys, numpy as np, torch
sys.path.insert(0, ".")
from bee_code_interpreter_fs_b200 import GpuHasher
h = GpuHasher(0)
# ragged batch: 151552 objects of 16..48 KiB, 16-byte aligned, longest first
rng = np.random.default_rng(5)
lengths = rng.integers(16384, 49152, size=151552).astype(np.int64)
off = np.concatenate([[0], np.cumsum((lengths + 15) // 16 * 16)[:-1]]).astype(np.int64)
buf = torch.empty(int(off[-1] + lengths[-1] + 64), dtype=torch.uint8, device="cuda:0")
h.gen_synthetic(buf, 3, 0, 1, buf.numel(), buf.numel())
d_off, d_len = torch.from_numpy(off).cuda(), torch.from_numpy(lengths).cuda()
order = torch.from_numpy(np.argsort(-lengths, kind="stable").astype(np.int32)).cuda()
for _ in range(3):
    h.digest_device(buf, d_off, d_len, order=order)
# few long objects: 592 x 4 MiB, one warp each
n, L = 592, 4 << 20
long_buf = torch.empty(n * L, dtype=torch.uint8, device="cuda:0")
h.gen_synthetic(long_buf, 4, 0, n, L)
for _ in range(3):
    h.digest_device_uniform(long_buf, L, L, n)
torch.cuda.synchronize()
print("ok")
PY
python /tmp/secondary.py > gpurun_out/ncu_secondary_plain_${TAG}.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"kernel_ldg|kernel_long" -c 6 -f -o gpurun_out/prof_secondary_${TAG} python /tmp/secondary.py > gpurun_out/ncu_secondary_${TAG}.log 2>&1
echo "secondary capture exit $?"
