import sys, numpy as np
sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tests')
from conftest import fixture_scene, CLEAR
from test_gpu_parity import _camera_for
from oracle.binding import OracleScene
from raytracer_b200 import librt
name='conference'; W=H=64
scene=fixture_scene(name); cam=_camera_for(name,1.0)
orc=OracleScene(scene).build_bvh()
t=orc.direct_reference(cam.ubo(),CLEAR,W,H,8)[...,:3].astype(np.float64)
lit=t.sum(axis=2)>1e-3
def stats(tag,g):
    rel=np.abs(g-t).sum(axis=2)[lit]/(t.sum(axis=2)[lit]+0.02)
    ratio=g.sum(axis=2)[lit]/t.sum(axis=2)[lit]
    print(tag,'rmse %.4f'%np.sqrt(((g-t)[lit]**2).mean()),'median rel err %.4f'%np.median(rel),'p90 %.4f'%np.percentile(rel,90),'max px %.2f'%g[lit].max(),'mean ratio %.4f'%(g[lit].mean()/t[lit].mean()),'median ratio %.4f'%np.median(ratio))
for flags in (0,1,2,3):
    c=librt.Context(device=0,radiance_clamp=1e30); c.build_scene(scene); c.resize(W,H)
    prm=librt.RtRestirParams(); c.L.rt_default_restir_params(prm); prm.flags=flags
    for f in range(256):
        c.restir_frame(cam.ubo(),CLEAR,f,prm)
        if f+1 in (8,64,256): stats('accum flags %d frames %3d'%(flags,f+1), c.readback()[...,:3].astype(np.float64))
    # single-frame quality with a warm history: every frame overwrites the image
    acc=[]
    for f in range(24):
        c.restir_frame(cam.ubo(),CLEAR,0 if False else 0,prm) if False else None
    c.close()
    c=librt.Context(device=0,radiance_clamp=1e30); c.build_scene(scene); c.resize(W,H)
    errs=[]
    for f in range(16):
        # distinct seeds per frame come from frame_number; overwrite needs frame 0, so read single frames via difference
        c.restir_frame(cam.ubo(),CLEAR,f,prm)
        img=c.readback()[...,:3].astype(np.float64)
        if f==0: prev=img.copy(); single=img
        else: single=img*(f+1)-prev*f; prev=img.copy()
        rel=np.abs(single-t).sum(axis=2)[lit]/(t.sum(axis=2)[lit]+0.02)
        errs.append(np.median(rel))
    print('single-frame median rel err, flags',flags,':',' '.join('%.3f'%e for e in errs))
    c.close()
