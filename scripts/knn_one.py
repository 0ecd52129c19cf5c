import sys; sys.path.insert(0, ".")
from scripts.knn_check import run
N, D, k = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
run(N, D, k, check=False); run(N, D, k, check=False)
