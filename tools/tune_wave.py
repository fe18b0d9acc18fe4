"""GPU experiment: bf_receive throughput vs wave size / streams in flight (BF_WAVE_LANES, BF_WAVE_STREAMS)."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from beifong_b200 import configs
s = configs.c2_car_body(1).commit()
n = 1 << 26
res = []
for lanes in (1 << 20, 1 << 21, 1 << 22, 1 << 23):
    for streams in (2, 4, 8, 16):
        os.environ["BF_WAVE_LANES"] = str(lanes); os.environ["BF_WAVE_STREAMS"] = str(streams)
        s.receive(n, seed=1); s.synchronize()
        ts = []
        for it in range(3):
            s.receive(n, seed=2 + it); s.synchronize(); ts.append(s.last_receive_ms()[0])
        res.append((lanes, streams, min(ts), n / min(ts) / 1e6))
        print(lanes, streams, "ms", min(ts), "Gsamples/s", n / min(ts) / 1e6, flush=True)
