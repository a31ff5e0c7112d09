import sys, json, ctypes, time
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/open-gpu-kernel-modules_b200")
import bench, uvmtier
lib = uvmtier.lib()
with uvmtier.Context(0) as ctx:
    t=time.time()
    out = bench.oversubscription_numbers(ctx, lib, uvmtier.P, 0)
    out["total_s"]=time.time()-t
print(json.dumps(out, indent=1))
