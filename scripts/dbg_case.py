import sys; sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tests')
import numpy as np, cases, oracle_binding as ob
import fractal_renderer_b200 as f
cfg = cases.all_small_cases()['mandel_huge_coords']
ref = ob.render_rows(cfg, threads=1)
it, z = f.render_counts(cfg)
rgb = f.get_image(cfg).reshape(cfg.height,cfg.width,3)
d = np.abs(rgb.astype(int)-ref['rgb'].astype(int)).max(-1)
ys,xs = np.nonzero(d)
print(len(ys), 'differ')
for y,x in list(zip(ys,xs))[:5]:
    print(x,y,'iters',it[y,x],ref['iters'][y,x],'z',z[y,x],ref['z'][y,x],'gpu',rgb[y,x],'cpu',ref['rgb'][y,x])
