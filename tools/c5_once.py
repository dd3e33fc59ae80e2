"""A reduced C5 pass for profiling: python tools/c5_once.py [tris=10000000] [rays=20000000]"""
import sys
sys.path.insert(0, ".")
import torch
from path_tracing_b200 import api, gen
tris = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
n = int(sys.argv[2]) if len(sys.argv) > 2 else 20_000_000
dev = torch.device("cuda", 0)
sc = api.Scene(gen.soup_scene(tris))
g = torch.Generator(device=dev); g.manual_seed(7)
org = torch.rand((n, 3), device=dev, generator=g)
d = torch.randn((n, 3), device=dev, generator=g); d = d / d.norm(dim=1, keepdim=True)
p2 = torch.rand((n, 3), device=dev, generator=g)
ids = torch.empty(n, dtype=torch.int32, device=dev); t = torch.empty(n, dtype=torch.float32, device=dev)
st = torch.cuda.current_stream().cuda_stream
for mode, second in ((api.CLOSEST, d), (api.ANY, p2), (api.CLOSEST | api.KEEP_RAY_ORDER, d), (api.ANY | api.KEEP_RAY_ORDER, p2)):
    for _ in range(2):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); sc.intersect_device(org.data_ptr(), second.data_ptr(), n, mode, ids.data_ptr(), t.data_ptr(), st); e1.record(); torch.cuda.synchronize()
    print("mode", mode, "Mrays/s", n / e0.elapsed_time(e1) / 1e3)
