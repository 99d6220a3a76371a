import sys
sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/direct-visual-odometry_b200'); sys.path.insert(0,'/root/repo/tests')
import numpy as np
from oracle.oracle import Oracle
from dvo_b200 import synth
import dvo_b200
from helpers import punch_invalid
orc=Oracle()
g, d, poses, K = synth.sequence(5, 640, 480, seed=77, trans_step=0.012, rot_step_deg=0.3)
g = np.stack([punch_invalid(x, seed=i, n_rect=2, frac_px=0.002) for i, x in enumerate(g)])
d0 = orc.cull_image(d[0], 2); s0 = np.full_like(d0, 0.3)
dv = dvo_b200.Dvo(K=K, max_frames=10, max_keyframes=10)
vo = orc.vo_create(K,3,2)
for i in range(3):
    # oracle track log: replicate by calling orc.track on copies before odometrize
    if i>0:
        ob = orc.frame_create(g[i], K, 3, 2)
        xo, lo = orc.track(ob, vo.ref_frame())
        for L in lo: print("   o",L["level"],L["iteration"],L["valid_pixels"],L["residual"],L["update"])
        if i==2:
            from helpers import oracle_gn_inputs
            kf=vo.ref_frame(); ks=dv.vo_ref_slot(0)
            gs=dv.frame_create(g[i])
            xi0=np.zeros(6,np.float32)
            upd,res,valid,mask=dv.gn_step(gs,ks,0,xi0)
            inp=oracle_gn_inputs(orc,kf,ob,0,xi0)
            o_upd,o_res,o_valid,o_mask=orc.optimize(*inp,xi0,0)
            print("gn valid",valid,o_valid,"maskdiff",int((mask[0]!=o_mask).sum()))
            ys,xs=np.nonzero(mask[0]!=o_mask)
            gd=dv.frame_download(ks,0,"depth")[0]; gg=dv.frame_download(gs,0,"gray")[0]
            gw=dv.warp_image(ks,0,xi0)[0]
            for y,x in list(zip(ys,xs))[:10]:
                print("  px",x,y,"g",mask[0][y,x],"o",o_mask[y,x],"depth",gd[y,x],inp[2][y,x],"I1",gg[y,x],inp[0][y,x],"I2",gw[y,x],inp[1][y,x])
    T_o = vo.odometrize(g[i], d0 if i==0 else None, s0 if i==0 else None)
    T_g = dv.vo_push(g[i], d0 if i==0 else None, s0 if i==0 else None)
    st = dv.vo_state()
    print("frame",i,"kf",st["was_keyframe"],vo.last_was_keyframe(),"rel",st["rel_xi"][0], vo.last_frame().xi()[1])
    for L in dv.vo_last_log()[0]:
        print("   g",L["level"],L["iteration"],L["valid_pixels"],L["residual"],L["update"])
    ks=dv.vo_ref_slot(0); kf=vo.ref_frame()
    for l in range(3):
        for kind in ("gray","gradx","grady","depth","sigma"):
            a=dv.frame_download(ks,l,kind)[0]; b=kf.image(l,kind)
            print("   ",l,kind,"bitexact",a.tobytes()==b.tobytes(), "ndiff",int((a!=b).sum()), "max",float(np.abs(a-b).max()))
    a=dv.frame_download(ks,2,"age")[0]; b=kf.image(2,"age"); print("    age",a.tobytes()==b.tobytes())
