import sys, math, numpy as np
sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tests')
import __graft_entry__ as ge
from test_gpu_tile import camera, setup
bre=ge.load_bre(); orc=ge.load_orc()
o=orc.Oracle(); ctx=bre.Bre(device=0)
n=40000
setup(o,ctx,n,32)
def diff(ro,rd,w,h,tag):
    ref_rgb, ref_offs, ref_idx, _ = o.query(ro, rd, float(n), want_hits=True, brute=True)
    ctx.set_ray_grid(w,h)
    rgb, offs, idx = ctx.query(ro, rd, float(n), want_hits=True)
    rc=np.diff(ref_offs.astype(np.int64)); gc=np.diff(offs.astype(np.int64))
    bad=np.nonzero(rc!=gc)[0]
    print(tag,"rays",len(rc),"bad rays",len(bad),"first bad",bad[:40].tolist(),"last bad",bad[-10:].tolist())
    print("  deficits", (rc-gc)[bad[:20]].tolist(), "ref", rc[bad[:20]].tolist())
    c=ctx.count(ro,rd); print("  counters",c)
    ctx.set_ray_grid(0,0)
ro,rd=o.synth_rays_incoherent(64*32)
diff(ro,rd,64,32,"incoh 64x32")
diff(ro,rd,64,32,"incoh 64x32 again")
ro,rd=camera(17,33,(0.5,0.5,-1.5),(0.5,0.5,0.5))
diff(ro,rd,17,33,"17x33")
ro,rd=camera(64,48,(0.5,0.5,-1.5),(0.5,0.5,0.5),jitter=3.0,seed=3)
diff(ro,rd,64,48,"jit3")
