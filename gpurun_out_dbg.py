import sys, os, time; sys.path.insert(0,'/root/repo')
import numpy as np
from beifong_b200 import configs
s=configs.c2_car_body(1).commit()
n=1<<23
for lanes in (1<<19, 1<<20, 1<<21):
    for streams in (1,2,4,8):
        os.environ["BF_WAVE_LANES"]=str(lanes); os.environ["BF_WAVE_STREAMS"]=str(streams)
        s.receive(n, seed=1); s.synchronize()
        ts=[]
        for it in range(3):
            s.receive(n, seed=2+it); s.synchronize(); ts.append(s.last_receive_ms()[0])
        print(lanes, streams, "ms", min(ts), "Gsamples/s", n/min(ts)/1e6)
