import sys, os
R=os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0]=[R, R+'/tests', R+'/oracle', R+'/tools']
import numpy as np
from conftest import make_scenarios, oracle_run
from oracle_py import Oracle
import humanoid_mujoco_b200 as hmj
o=Oracle()
scen=make_scenarios(o,48,seed=11,settle=(0,500))
qpos,qvel,ws,ctrl,params=scen
b=hmj.HumanoidBatch(48); b.set_params(params); b.set_state(qpos,qvel,np.zeros(48),ws); b.set_ctrl(ctrl)
b.forward()
qacc,qfc,nefc,niter,qas=b.get('qacc'),b.get('qfrc_constraint'),b.get('nefc'),b.get('niter'),b.get('qacc_smooth')
def rel(a,b): return float(np.max(np.abs(a-b))/max(np.max(np.abs(b)),1.0))
for i in range(48):
    r=oracle_run(o,*[s[i] for s in scen],0,fn='forward')
    e1,e2,e3=rel(qacc[i],r['qacc']),rel(qfc[i],r['qfrc_constraint']),rel(qas[i],r['qacc_smooth'])
    flag = '  <<<<' if max(e1,e2)>1e-9 else ''
    print(i,'nefc',nefc[i],int(r['nefc']),'ncon',int(r['ncon']),'niter',niter[i],int(r['solver_niter']),'nls',int(r['solver_nls']),'err qacc %.2e qfc %.2e qas %.2e'%(e1,e2,e3),flag)
