import os, sys, json
ROOT='/root/repo'
sys.path[:0]=[ROOT, ROOT+'/dex-ct-sim_b200', ROOT+'/tests']
import numpy as np, torch
from dexct import synthetic as S, config
from xtomosim.system import FanBeamGeometry, VoxelPhantom, xRaySpectrum
from xtomosim.forward_project import get_sino
import matdecomp
from oracle import gn_oracle
G=ROOT+'/tests/golden'
n=128
ph=VoxelPhantom('p',None,None,n,n,1,dx=51.2/n,dy=51.2/n,dz=0.1,labels=S.phantom_labels(n),materials=S.MATERIALS)
ct=FanBeamGeometry(200,180,0.8230337,60.0,100.0,1.0,True,G+'/input/detector/eta_eid_mv.bin',rotation=6.283185)
specs=[]
sinos=[]
for sid,d in (('detunedMV',9),('80kV',1)):
    sp=xRaySpectrum(G+f'/input/spectrum/{sid}_1mGy_float32.bin',sid); sp.rescale_counts(ct.A_iso*d/ct.N_proj)
    specs.append(sp); sinos.append(get_sino(ct,ph,sp)[0])
from dexct.ops import gn_decompose
ee,i0,mus=gn_oracle.gn_model_tables(ct,specs[0],specs[1],matdecomp.xc.mixatten)
r32=lambda x: np.asarray(x).astype(np.float32).astype(np.float64)
g=r32(np.stack(sinos)); i0=r32(i0); mus=r32(mus)
for dt in ('f32','f64'):
    out=gn_decompose(g,i0,mus,50,dtype=dt).cpu().numpy().astype(np.float64)
    m1,m2=out[...,0],out[...,1]
    bad=~(np.isfinite(m1)&np.isfinite(m2))
    print(dt,'non-finite pixels',bad.sum(),'of',bad.size, 'max', np.nanmax(np.abs(m1)), np.nanmax(np.abs(m2)))
    if dt=='f32': bad32=bad; m32=(m1,m2)
    else: m64=(m1,m2)
vv=np.unique(np.where(bad32)[0])[:3]
print('views with bad pixels', vv)
for v in vv:
    nB=g.shape[2]
    i0_t=np.tile(i0,(1,nB)).reshape(2,nB,len(ee))
    with np.errstate(all='ignore'):
        ref=gn_oracle.optimize_sino_cpu(g[:,v:v+1],ee,i0_t,mus,50)
    cols=np.where(bad32[v])[0]
    print('view',v,'bad cols',cols[:10])
    for c in cols[:5]:
        print('  col',c,'g',g[:,v,c],'oracle',ref[0,c],'gpu64',m64[0][v,c],m64[1][v,c],'gpu32',m32[0][v,c],m32[1][v,c])
ok=~bad32
d=np.maximum(np.abs(m32[0]-m64[0]),np.abs(m32[1]-m64[1])); sc=np.maximum(np.abs(m64[0]),np.abs(m64[1]))+1.0
print('pixels with |f32-f64| > 1e-3 (1+|a|):', int((d>1e-3*sc).sum()), ' > 1e-4:', int((d>1e-4*sc).sum()), 'of', d.size)
print('f32 vs f64 on finite pixels: ', np.max(np.abs(m32[0][ok]-m64[0][ok])), np.max(np.abs(m64[0][ok])))
idx=np.argwhere(d>1e-4*sc)
for v,c in idx[:12]:
    print('pixel',v,c,'g',g[:,v,c],'f32',m32[0][v,c],m32[1][v,c],'f64',m64[0][v,c],m64[1][v,c])
