import os, sys
sys.path.insert(0, "/root/repo")
from beifong_b200 import configs
s = configs.c2_car_body(1).commit()
n = 1 << 28
for lanes in (1 << 23, 1 << 24, 1 << 25):
    for streams in (4, 8):
        os.environ["BF_WAVE_LANES"] = str(lanes); os.environ["BF_WAVE_STREAMS"] = str(streams)
        s.receive(n, seed=1); s.synchronize()
        ts = []
        for it in range(2):
            s.receive(n, seed=2 + it); s.synchronize(); ts.append(s.last_receive_ms()[0])
        print(lanes, streams, "ms", min(ts), "Gsamples/s", n / min(ts) / 1e6, flush=True)
