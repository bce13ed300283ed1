import os, sys, json
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import bench_configs as bc
import torch
for rec in bc.config3(torch, bc.chain_of("panda"), ticks=6):
    print(json.dumps(rec))
