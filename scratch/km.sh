cd /root/repo
timeout 600 python -m pytest tests/test_gpu_cluster.py -m gpu -q 2>&1 | tail -8
for k in 16 32; do timeout 300 python profiles/prof_kmeans.py 50000000 $k 2>&1 | tail -1; done
