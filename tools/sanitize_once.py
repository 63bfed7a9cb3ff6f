"""One pass over every kernel family at small shapes — the target of `compute-sanitizer --tool memcheck|racecheck|synccheck`.
    compute-sanitizer --tool memcheck python tools/sanitize_once.py
Shapes are tiny (the tools slow kernels down 20-100x) but walk the same code paths as the benchmarked ones: dim=48 puts
C = 48 / 96 on the fused LN+1x1+dw+gate kernel and C = 192 / 384 on the two-kernel path; bf16 = tcgen05 / TMA kernels, fp32 =
CUDA-core kernels; one training step runs every backward kernel and AdamW; a small MoCE-IR runs the expert / router kernels."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import b200ir
from oracle import restormer_oracle as O

torch.manual_seed(0)
cfg = dict(O.DEFAULT_CFG)
cfg.update(dim=48, num_blocks=[1, 1, 1, 1], num_refinement_blocks=1)
sd = O.synth_state_dict(dict(dim=48, num_blocks=[1, 1, 1, 1], num_refinement_blocks=1), seed=3)
img = O.synth_image(2, 64, 64, seed=7).cuda()
net = b200ir.Restormer(**cfg).cuda().eval()
net.load_state_dict(sd, strict=True)
with torch.no_grad():
    y32 = net(img)
    y16 = net.set_compute_dtype(torch.bfloat16)(img)
torch.cuda.synchronize()
print("forward fp32/bf16 ok", float((y32 - y16).abs().max()))
net.train()
opt = b200ir.AdamW(net.parameters(), lr=2e-4)
clean = O.synth_image(2, 64, 64, seed=8).cuda()
for dt in (torch.bfloat16, torch.float32):
    net.set_compute_dtype(dt)
    loss = b200ir.train_step(net, opt, img, clean)
    torch.cuda.synchronize()
    print("train step", dt, float(loss))
try:
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle"))
    from gen_golden_cases import MOCEIR_SMALL
    m = b200ir.MoCEIR(**MOCEIR_SMALL).cuda().eval()
    with torch.no_grad():
        for dt in (torch.float32, torch.bfloat16):
            m.set_compute_dtype(dt)
            ym = m(O.synth_image(2, 64, 64, seed=9).cuda())
    torch.cuda.synchronize()
    print("moce-ir forward ok", tuple(ym.shape))
except ImportError as e:
    print("moce-ir part skipped:", e)
from b200ir import pipeline  # noqa: E402,F401
a, b = torch.rand(2, 3, 48, 56, device="cuda"), torch.rand(2, 3, 48, 56, device="cuda")
print("psnr/ssim", b200ir.psnr_ssim(a, b).cpu().tolist())
print("sanitize_once done")
