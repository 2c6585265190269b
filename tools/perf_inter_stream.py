"""Development probe: vvcb200_inter_batch on the default stream and on a stream of its own (the forked width classes must
overlap either way)."""
import sys, os
sys.path.insert(0,'/root/repo')
import numpy as np, torch
import bench
from ffmpeg_b200 import lib, hostutil as util
pic=bench.make_picture(2026,3840,2160,0.1); it=pic["inter"]
dev=lambda a: torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1)).cuda()
refs=[util.to_device_planes(r,torch) for r in it["refs"]]; refs_p=[lib.device_planes(r) for r in refs]
cur=util.to_device_planes([np.zeros((h,w),np.uint16) for (w,h) in pic["dims"]],torch); cur_p=lib.device_planes(cur)
blocks=dev(it["blocks"]); prof=dev(it["prof"])
def run(stream):
    q=None if stream is None else stream.cuda_stream
    for _ in range(5): lib.inter_batch(cur_p,refs_p,blocks,len(it["blocks"]),it["params"],prof=prof,stream=q)
    torch.cuda.synchronize()
    e0,e1=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
    s=torch.cuda.current_stream() if stream is None else stream
    e0.record(s)
    for _ in range(30): lib.inter_batch(cur_p,refs_p,blocks,len(it["blocks"]),it["params"],prof=prof,stream=q)
    e1.record(s); torch.cuda.synchronize()
    return e0.elapsed_time(e1)/30
print("default stream %.4f ms"%run(None))
print("explicit stream %.4f ms"%run(torch.cuda.Stream()))
print("default stream %.4f ms"%run(None))
