"""One launch (after a warm-up launch) of every hot kernel at its BASELINE shape, for `ncu --set full`:
    python scripts/prof_kernels.py
Kernels: conv_halo_kernel<64> resident 3x3 64->64 (+res) @1088x1920, streamed 3x3 128->128 @272x480 and
192->192 @68x120, 7x7 32->64 @1088x1920, conv_tc_kernel 3x3 s2 64->64 @1088x1920, the split-bf16 (parity) 3x3 64->64
@544x960 with its split kernel, se partial/finish, dual_prior_pass0, flow_warp8_bf16 64 ch @1088x1920."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from vcm_ts_b200.backend import CudaBackend, View  # noqa: E402
from vcm_ts_b200.entropy import ScaleIndexTable  # noqa: E402

bf = torch.bfloat16
dev = "cuda"
B = CudaBackend("cuda:0", "bf16")
BP = CudaBackend("cuda:0", "parity")
g = torch.Generator().manual_seed(0)


def act(C, H, W, dt=bf):
    return View(torch.randn((1, H, W, C), device=dev).to(dt))


def conv(Bk, K, stride, Cin, Cout, H, W, res=False, dt=bf):
    w = torch.randn((Cout, Cin, K, K), generator=g) / (Cin * K * K) ** 0.5
    pc = Bk.pack_conv(w.cuda(), torch.zeros(Cout).cuda())
    x = act(Cin, H, W, dt)
    r = act(Cout, H // stride, W // stride, dt) if res else None
    for _ in range(2):
        Bk.conv(x, pc, stride=stride, post=0.01, res=r)
    torch.cuda.synchronize()


conv(B, 3, 1, 64, 64, 1088, 1920, res=True)        # conv_rows_kernel<64, 2> (2 output rows stacked along N)
os.environ["VCM_TC_NO_ROWS"] = "1"
conv(B, 3, 1, 64, 64, 1088, 1920, res=True)        # the same launch on conv_halo_kernel<64>
os.environ.pop("VCM_TC_NO_ROWS")
conv(B, 3, 1, 64, 32, 1088, 1920, res=True)        # conv_rows_kernel<64, 4>
conv(B, 1, 1, 64, 64, 1088, 1920, res=True)        # 1x1 + residual: HBM-bound
conv(B, 3, 1, 64, 64, 544, 960, res=True)
conv(B, 3, 1, 128, 128, 272, 480, res=True)
conv(B, 3, 1, 192, 192, 68, 120)
conv(B, 7, 1, 32, 64, 1088, 1920)
conv(B, 3, 2, 64, 64, 1088, 1920)
conv(BP, 3, 1, 64, 64, 544, 960, res=True, dt=torch.float32)
# SE layer (mean + FC), 64 ch at 544x960
x = act(64, 544, 960)
w1 = torch.randn((4, 64), device=dev) / 8
w2 = torch.randn((64, 4), device=dev)
for _ in range(2):
    B.se_scale(x, w1, w2)
# dual-prior pass 0 (encode) at the 1080p latent size
C, h, w = 96, 68, 120
table = ScaleIndexTable("laplace")
fusion, y = act(3 * C, h, w), act(C, h, w)
params = B.new(1, h, w, 4 * C)
yq = B.empty((1, h, w, C), torch.float32)
sym, idx = B.empty((1, C // 2, h, w), torch.int16), B.empty((1, C // 2, h, w), torch.uint8)
for _ in range(2):
    B.dp_pass0(0, table, fusion, params, C, y=y, yq=yq, sym=sym, idx=idx)
# flow_warp 64 ch bf16 at 1088x1920
feat = act(64, 1088, 1920)
flow = View((torch.randn((1, 1088, 1920, 2), device=dev) * 3).contiguous())
for _ in range(2):
    B.warp(feat, flow)
torch.cuda.synchronize()
print("ok")
