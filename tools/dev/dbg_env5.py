import sys, os
R=os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0]=[R, R+'/tests', R+'/oracle', R+'/tools']
import numpy as np
from conftest import make_scenarios, oracle_run
from oracle_py import Oracle
import humanoid_mujoco_b200 as hmj
E=int(sys.argv[1]) if len(sys.argv)>1 else 5
o=Oracle()
scen=make_scenarios(o,48,seed=11,settle=(0,500))
qpos,qvel,ws,ctrl,params=scen
print("=== oracle forward env",E, file=sys.stderr, flush=True)
r=oracle_run(o,*[s[E] for s in scen],0,fn='forward')
print("=== gpu", flush=True)
b=hmj.HumanoidBatch(48); b.set_params(params); b.set_state(qpos,qvel,np.zeros(48),ws); b.set_ctrl(ctrl)
b.forward()
