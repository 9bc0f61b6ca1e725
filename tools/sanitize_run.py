"""Every entry point of libmerl_b200 once at small sizes (for compute-sanitizer memcheck / racecheck).
Sizes are chosen to hit the vector path, the scalar tail, the redo queue (edge-stress queries) and
a multi-pass, multi-table render."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ctypes as C
import numpy as np, torch
from mitsuba_customization_b200 import _cabi, bsdf, render as RD, synth

plug = bsdf.MERL(data=synth.synthetic_table(0))
ctx = bsdf.BSDFContext()
for n, maker in ((4099, lambda n: synth.random_queries(n, seed=1)), (1 << 14, lambda n: synth.edge_stress_queries(n))):
    wi, wo = maker(n)
    n = wi.shape[1]
    si, two = bsdf.SurfaceInteraction3f(wi=torch.from_numpy(wi).cuda()), torch.from_numpy(wo).cuda()
    plug.bins(si, two); plug.eval(ctx, si, two); plug.pdf(ctx, si, two); plug.eval_pdf(ctx, si, two)
    u1, u2 = torch.rand(n, device="cuda"), torch.rand(2, n, device="cuda")
    for mode in (0, 1):
        _cabi.lib().merl_set_sample_search(mode)
        plug.sample(ctx, si, u1, u2)
    _cabi.lib().merl_set_sample_search(0)
    g = torch.randn(3, n, device="cuda")
    for agg in (True, False):
        plug.backward(si, two, g, warp_aggregate=agg)
    wi4 = torch.zeros(n, 4, device="cuda"); wi4[:, :3] = si.wi.T
    wo4 = torch.zeros(n, 4, device="cuda"); wo4[:, :3] = two.T
    out4, o2 = torch.empty(n, 4, device="cuda"), torch.empty(n, 4, device="cuda")
    s = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    _cabi.check(_cabi.lib().merl_eval_pdf_aos(plug._h, wi4.data_ptr(), wo4.data_ptr(), n, out4.data_ptr(), s))
    u4 = torch.rand(n, 4, device="cuda")
    _cabi.check(_cabi.lib().merl_sample_aos(plug._h, wi4.data_ptr(), u4.data_ptr(), n, out4.data_ptr(), o2.data_ptr(), s))
# big-batch backward path (n >= cells) on a small customized table
dims = (16, 16, 32)
small = bsdf.CustomizedMeasurement(data=synth.synthetic_table(2, dims=dims, theta_h_sqrt=False), resolution=dims,
                                   theta_h_mapping="linear")
wi, wo = synth.random_queries(3 * 16 * 16 * 32 + 1, seed=3)
small.backward(bsdf.SurfaceInteraction3f(wi=torch.from_numpy(wi).cuda()), torch.from_numpy(wo).cuda(),
               torch.ones(3, wi.shape[1], device="cuda"))
# host-buffer entry point
hw, ho = torch.from_numpy(wi).pin_memory(), torch.from_numpy(wo).pin_memory()
small.eval_pdf_host(hw, ho, torch.empty(4, wi.shape[1]).pin_memory())
# render: two tables, several passes, graph replay and plain launches
sc = RD.grid_scene([plug, bsdf.MERL(data=synth.synthetic_table(1))], 2, 2)
for graphs in (1, 0):
    _cabi.lib().merl_set_render_graphs(graphs)
    RD.render(sc, 48, 32, 4, max_depth=3, seed=2, wavefront=2048)
_cabi.lib().merl_set_render_graphs(1)
torch.cuda.synchronize()
print("sanitize_run ok,", _cabi.launch_count(), "launches")
