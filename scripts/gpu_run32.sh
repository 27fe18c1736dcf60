timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 300 python profiles/nearest_bench.py 4096 2>&1 | tail -1
