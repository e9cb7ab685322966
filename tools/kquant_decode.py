import sys, os, json
sys.path.insert(0, '/root/repo'); 
import bench, numpy as np
sys.path.insert(0, os.path.join('/root/repo', 'ggml-hexagon_b200'))
import b200ops as B
s = B.Session(0)
print(json.dumps({k: v for k, v in bench.time_kquant(B, s, 10).items() if k != 'pp512'}))
