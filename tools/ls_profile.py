import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from routest_b200 import Engine
from tools.dist_search_check import build
eng = Engine(0)
stops = int(sys.argv[1]) if len(sys.argv) > 1 else 10000
pr, trips = build(eng, 5678, stops, stops // 100)
batched = len(sys.argv) > 2 and sys.argv[2] == "batched"
moves, cost, _ = pr.local_search(max_moves=16 if batched else 64, batched=batched)
print(moves, cost)
