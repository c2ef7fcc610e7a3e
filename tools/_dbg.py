import sys; sys.path.insert(0,'.')
import numpy as np, oracle
from serverless_learn_b200 import Replica, SlError
orc=oracle.Oracle()
def run(n,F,B,steps,grid=0,variant=0):
    X,y=orc.synth(0,n,F)
    w=np.zeros(F); lo=orc.lr_steps(X,y,B,0,steps,w,0.1)
    with Replica(F,B,eta=0.1) as r:
        r.set_option("path",1); r.set_option("grid",grid); r.set_option("variant",variant)
        r.load_shard(X,y)
        try:
            lg=r.train_steps(steps)
            print(n,F,B,steps,grid,variant,"ok werr",np.abs(r.model()-w).max()/np.abs(w).max(), "slots",r.info("ring_slots"))
        except SlError as e: print(n,F,B,steps,grid,variant,"ERR",e)
run(64,8,16,6); run(64,8,16,1); run(64,8,16,2); run(64,8,16,6,grid=1); run(64,8,16,6,grid=4); run(64,8,16,6,grid=16)
run(640,8,160,6); run(64,128,16,6); run(64,1024,16,6); run(4096,1024,148,6); run(4096,1024,148,6,variant=2); run(64,8,16,6,variant=2)
