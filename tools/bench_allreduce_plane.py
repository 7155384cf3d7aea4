"""Data-plane comparison of the butterfly all-reduce between two peers on localhost: the C++ thread pipeline
(native/allreduce_pipe.cpp) against the asyncio + msgpack part loop, same matchmaking, same codecs.
    python tools/bench_allreduce_plane.py [million_floats=32]"""
import json
import os
import sys
import threading
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from dalle_b200.collab import compression as C
from dalle_b200.collab.averager import DecentralizedAverager
from dalle_b200.dht import DHT

N = int(float(sys.argv[1]) * 1e6) if len(sys.argv) > 1 else 32_000_000


def run(plane: str, codec: str):
    os.environ["DALLE_B200_NATIVE_ALLREDUCE"] = "1" if plane == "native" else "0"
    comp = {"none": C.NoCompression(), "fp16": C.Float16Compression(),
            "adaptive": C.SizeAdaptiveCompression(2 ** 16 + 1, C.Float16Compression(), C.Uniform8BitQuantization())}[codec]
    first = DHT(start=True, host_maddrs=["/ip4/127.0.0.1/tcp/0"])
    dhts = [first, DHT(initial_peers=[str(m) for m in first.get_visible_maddrs()], start=True,
                       host_maddrs=["/ip4/127.0.0.1/tcp/0"])]
    try:
        g = torch.Generator().manual_seed(0)
        avs = [DecentralizedAverager([torch.randn(N, generator=g)], dhts[i], prefix=f"bp_{plane}_{codec}",
                                     min_matchmaking_time=1.0, target_group_size=2, compression=comp,
                                     allreduce_timeout=300, sender_timeout=60) for i in range(2)]
        best = 1e9
        for _ in range(2):
            times = [0.0, 0.0]

            def work(i):
                t0 = time.time()
                avs[i].step(weight=1.0, timeout=300)
                times[i] = time.time() - t0

            ts = [threading.Thread(target=work, args=(i,)) for i in range(2)]
            [t.start() for t in ts]
            [t.join() for t in ts]
            best = min(best, max(times))
        assert all(a.native_rounds == (2 if plane == "native" else 0) for a in avs)
        return best
    finally:
        [d.shutdown() for d in dhts]


out = {"floats": N, "peers": 2, "host_cores": os.cpu_count()}
for codec in ("none", "adaptive"):
    for plane in ("asyncio", "native"):
        t = run(plane, codec)
        out[f"{codec}_{plane}_s"] = round(t, 3)
        print(codec, plane, f"{t:.3f} s per round (incl. ~1 s matchmaking window)", flush=True)
print(json.dumps(out))
