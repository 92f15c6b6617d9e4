for s in 1 2 3; do GNNB_KNN_SPLITS=$s timeout 120 python scripts/perf_knn_fp4.py 15625 50000 32 2>&1 | tail -1; done
for s in 1 2 8; do GNNB_KNN_SPLITS=$s timeout 200 python scripts/perf_knn_fp4.py 90000 50000 32 2>&1 | tail -1; done
for s in 1 2 4 8; do GNNB_KNN_SPLITS=$s timeout 120 python scripts/perf_knn_fp4.py 3534 52843 10 2>&1 | tail -1; done
for s in 2 4 8 16; do GNNB_KNN_SPLITS=$s timeout 120 python scripts/perf_knn_fp4.py 1814 10346 5 2>&1 | tail -1; done
