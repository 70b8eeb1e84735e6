import json, os, sys, subprocess
sys.path.insert(0, '/root/repo')
os.chdir('/root/repo')
import bench, torch
from panvc3_b200 import build as b; b.build()
from panvc3_b200.projector import MsaIndex
from panvc3_b200.workload import sorted_bam_window
dev = torch.device('cuda', 0); torch.cuda.set_device(0)
peak, _ = bench.measured_peak_gbs()
wl = bench.load_workload(1_000_000, 0, 'chr26')
idx = MsaIndex.from_rows(*wl.index_inputs, device=0)
c = idx.ctx; c.set_params(0, -1)
stream = torch.cuda.Stream(device=dev); c.set_stream(stream.cuda_stream)
sb, sm = sorted_bam_window(wl.msa, 1_000_000, bench.READ_LEN, bench.SEED + 5)
res = {}
for v in ('0', '2', '4', '3'):
    os.environ['PV3_K1T_VARIANT'] = v
    with torch.cuda.stream(stream):
        a = bench.device_leg(c, dev, stream, wl, wl.batch, wl.mates, 5, 3, peak)
        s = bench.device_leg(c, dev, stream, wl, sb, sm, 5, 3, peak)
    res[v] = {'random': [a['ms_per_step'], a['roofline']['kernel_ms']], 'sorted_window': [s['ms_per_step'], s['roofline']['kernel_ms']]}
    print(v, res[v], flush=True)
json.dump(res, open('gpurun_out/r2_v25_chr_layout_variants.json', 'w'))
