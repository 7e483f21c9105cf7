import sys, os, json, torch
ROOT='/root/repo'
sys.path[:0]=[ROOT, os.path.join(ROOT,'structured-gaussian-splatting_b200'), os.path.join(ROOT,'tests')]
from gsr_b200 import synth, dp
import diff_gaussian_rasterization as dgr
from diff_gaussian_rasterization.batched import BatchedRasterizationSettings, GaussianRasterizerBatched
spec=synth.CONFIGS["C4"]; dev=torch.device("cuda",0); V=8; S=16
scenes=[dp.make_leaf_params(synth.spec_scene(spec, seed=s), dev) for s in range(S)]
cams=[synth.make_camera(k, V, spec.W, spec.H, spec.radius, spec.fovx_deg).to(dev) for k in range(V)]
bg=synth.spec_background(spec, dev); go=synth.grad_image(spec.W, spec.H, device=dev)
vm=torch.stack([c.viewmatrix for c in cams]); pm=torch.stack([c.projmatrix for c in cams]); cp=torch.stack([c.campos for c in cams])
st={k: torch.stack([p[k].detach() for p in scenes]).requires_grad_(True) for k in scenes[0]}
e=lambda t,*d: t.unsqueeze(0).expand(S,*d).contiguous()
rs=BatchedRasterizationSettings(spec.H, spec.W, cams[0].tanfovx, cams[0].tanfovy, bg, 1.0, e(vm,-1,-1,-1), e(pm,-1,-1,-1), spec.D, e(cp,-1,-1), False)
gos=go[None,None].expand(S,V,-1,-1,-1).contiguous()
C=dgr._C
def run():
    m2=torch.zeros(S,V,spec.P,3,device=dev,requires_grad=True)
    c,r=GaussianRasterizerBatched(rs)(st["means3D"],m2,st["opacities"],st["shs"],st["scales"],st["rotations"])
    c.backward(gos)
    for t in st.values(): t.grad=None
    return r
for _ in range(2): r=run()
C.timing_reset(); C.timing_enable(True)
torch.cuda.synchronize()
e0=torch.cuda.Event(enable_timing=True); e1=torch.cuda.Event(enable_timing=True); e0.record()
n=4
for _ in range(n): run()
e1.record(); torch.cuda.synchronize()
C.timing_enable(False)
print("ms per call", e0.elapsed_time(e1)/n, "per view", e0.elapsed_time(e1)/n/(S*V), "visible frac", float((r>0).float().mean()))
print({s_: round(ms/max(c_,1),4) for s_,ms,c_ in C.timing_collect() if c_})
