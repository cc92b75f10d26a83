import sys, os
R=os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0]=[R, R+'/tests', R+'/oracle', R+'/tools']
import numpy as np
from conftest import make_scenarios
from oracle_py import Oracle, OracleBatch
import humanoid_mujoco_b200 as hmj
N=int(sys.argv[1]) if len(sys.argv)>1 else 256
o=Oracle()
scen=make_scenarios(o,N,seed=11,settle=(0,500))
qpos,qvel,ws,ctrl,params=scen
b=hmj.HumanoidBatch(N); b.set_params(params); b.set_state(qpos,qvel,np.zeros(N),ws); b.set_ctrl(ctrl)
ob=OracleBatch(N); ob.params[:]=params; ob.set_state(qpos,qvel,ws); ob.data['ctrl'][:]=ctrl
def rel(a,b): return np.max(np.abs(a-b),axis=1)/np.maximum(np.max(np.abs(b),axis=1),1.0)
tot_con=0
for step in range(1,11):
    b.step(None,1); ob.step(None,1,8)
    q,v,_,w=b.get_state()
    e=np.maximum(rel(q,ob.data['qpos']),rel(v,ob.data['qvel']))
    nc=b.get('ncon'); tot_con+=int(ob.data['ncon'].sum())
    print('step',step,'ncon mismatch',int((nc!=ob.data['ncon']).sum()),'envs>1e-9:',int((e>1e-9).sum()),'envs>1e-6:',int((e>1e-6).sum()),'max',e.max(),'contacts so far',tot_con, flush=True)
