import sys; sys.path.insert(0,'.')
import numpy as np
from loadbalancedchemistrymodel_b200 import capi, workloads as W
from oracle import oracle as O
mech,Y,T,p,rho=W.ch4_mixing_line(4000)
print(mech.species, mech.names, mech.lhs_exp, mech.rhs_exp, mech.arr)
opts=capi.default_options(solver="EulerImplicit",abs_tol=1e-12,rel_tol=1e-2,cTauChem=1.0,eq_rate_limiter=0)
chem=capi.Chemistry(mech,opts)
for dt in (1e-6,1e-9):
    got=chem.solve_host(Y,T,p,rho,dt,1e-7)
    ref=O.OracleModel(mech,opts).solve_batch(Y,T,p,rho,dt,1e-7,nthreads=8)
    e=np.abs(got["deltaTChem"]-ref["deltaTChem"])/ref["deltaTChem"]
    i=int(np.argmax(e))
    print("dt",dt,"worst",i,e[i],"cost",got["cost"][i],ref["attempts"][i],"T",T[i],"dtc",got["deltaTChem"][i],ref["deltaTChem"][i],"Tend diff",got["T"][i]-ref["T"][i], "n>1e-9", (e>1e-9).sum())
got=chem.solve_host(Y,T,p,rho,1e-6,1e-7); ref=O.OracleModel(mech,opts).solve_batch(Y,T,p,rho,1e-6,1e-7,nthreads=8)
sc=np.abs(ref["RR"]).max(0,keepdims=True); e=np.abs(got["RR"]-ref["RR"])/(sc+1e-30)
s,i=np.unravel_index(np.argmax(e),e.shape)
print("worst RR species",mech.species[s],"cell",i,"T",T[i],"Y",Y[:,i],"cost",got["cost"][i])
print("RR gpu",got["RR"][:,i]); print("RR ref",ref["RR"][:,i]); print("Tend",got["T"][i],ref["T"][i],"dtc",got["deltaTChem"][i],ref["deltaTChem"][i])
print("cells e>1e-6:",(e.max(0)>1e-6).sum(), "T of those", np.sort(T[e.max(0)>1e-6])[:10])
