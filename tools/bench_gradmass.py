#!/usr/bin/env python3
"""Gradient mass matrices (compute_gradient_mass_matrix) of bilinear quads: Kronecker kernel (opt-in, ib200_ctx_set_kron_gradient)
against the default path (sum-factorised kernels; for matching spaces followed by the in-place mirror pass that reproduces the
reference's lower-triangle mirroring -- round 2 ran those on the generic kernel).
Device resident, CUDA events, best of 3.   python tools/bench_gradmass.py"""
import sys, json; sys.path.insert(0,'/root/repo')
import numpy as np, interplib_b200 as ib
from interplib_b200 import DEVICE, synthetic as syn
ctx=ib.Context(0)
for p,E in ((4,1<<18),(8,1<<14),(16,1<<11)):
    basis,rules,gb=[("lagrange-gauss-lobatto",p)]*2,[("gauss",p+1)]*2,[("lagrange-uniform",1)]*2
    lower=[("lagrange-gauss-lobatto",p-1),("lagrange-gauss-lobatto",p)]
    N=(p+1)**2
    side=int(np.ceil(np.sqrt(E))); geom=np.ascontiguousarray(syn.quad_mesh_geometry(side*side,seed=1).reshape(-1,2,4)[:E])
    d_g=ctx.alloc(geom.nbytes).upload(geom); d_M=ctx.alloc(E*N*N*8)
    for name,bo in (("matching",basis),("mixed",lower)):
        No=int(np.prod([o+1 for _,o in bo]))
        for kron in (True,False):
            ctx.set_kron_gradient(kron); best=1e30
            for i in range(4):
                ctx.timer_start(2); ctx.mass_matrix_mapped(basis,bo,rules,gb,d_g.ptr,2,idx_in=1,idx_out=0,out=d_M.ptr,where=DEVICE,n_elements=E); ctx.timer_stop(2)
                ms=ctx.timer_ms(2); best=min(best,ms) if i else best
            ctx.set_kron_gradient(False)
            print(json.dumps({"p":p,"spaces":name,"kernel":"kronecker (opt-in)" if kron else "default","elements":E,"ms":round(best,3),"el_per_s":E/(best*1e-3),"GB/s":round(8*No*N*E/(best*1e-3)/1e9,1)}),flush=True)
