"""cfg3 (SDF mode): would a hierarchical cull pay?  For the sphere samples of one scored iteration on the cfg3 scene (CPU, replica
oracle): the share whose cell-level lower bound (min of the 8 corners) cannot rule out a contact, the share in contact, and
the 32-byte sectors a point would touch in an FP16 cell-minimum grid.  Output committed as profiles/r02_sdf_cull_stats.txt."""
import sys, numpy as np
sys.path.insert(0,'/root/repo')
from tests.helpers import sdf_case, o2_for, SEED
from tests.test_oracle_agreement import _host_state
case = sdf_case(n_samples=510, check_points=64, n_points=5, dims=(128,128,128))
o2 = o2_for(case)
nodes = o2.sdf_nodes()   # [nz][ny][nx]
sp = o2.spheres()
psi_col, psi_arc, base, mean, sigma = _host_state(case)
r = o2.iterate(psi_col, None, base, mean, sigma, np.zeros_like(mean), 0, SEED, 0, want_points=True)
pts = r['pts'].reshape(-1,4)   # x,y,z,yaw
print('n_success', r['n_success'], 'of', len(r['status']))
lo = np.array([-0.8,-0.8,0.0]); hi=np.array([0.8,0.8,0.9]); dims=np.array([128,128,128]); h=(hi-lo)/(dims-1)
c,s = np.cos(pts[:,3]), np.sin(pts[:,3])
tot=near=contact=0; sect=[]
cellmin = np.minimum.reduce([nodes[k:k+127, j:j+127, i:i+127] for k in (0,1) for j in (0,1) for i in (0,1)])
for a in range(0,len(pts),7):
    R = np.array([[c[a],-s[a],0],[s[a],c[a],0],[0,0,1]])
    X = pts[a,:3] + sp[:,:3]@R.T
    g = np.clip((X-lo)/h, 0, dims-1); ijk = np.minimum(g.astype(int), dims-2)
    cm = cellmin[ijk[:,2], ijk[:,1], ijk[:,0]]
    f = g-ijk
    # trilinear
    v = np.zeros(len(X))
    for dz in (0,1):
      for dy in (0,1):
        for dx in (0,1):
          w = (f[:,0] if dx else 1-f[:,0])*(f[:,1] if dy else 1-f[:,1])*(f[:,2] if dz else 1-f[:,2])
          v += w*nodes[ijk[:,2]+dz, ijk[:,1]+dy, ijk[:,0]+dx]
    tot += len(X); near += (cm - sp[:,3] < 1e-5).sum(); contact += (v - sp[:,3] < 0).sum()
    sect.append(len(set(zip(ijk[:,2], ijk[:,1], ijk[:,0]//16))))
print('samples', tot, 'near(cellmin)', near/tot, 'contact', contact/tot, 'fp16 sectors per point', np.mean(sect))
