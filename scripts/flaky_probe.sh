for i in 1 2 3 4 5 6; do python -m pytest tests/test_gpu_optimizer.py -m gpu -q 2>&1 | tail -1; done
echo PDL0
for i in 1 2 3 4 5 6; do NGP_PDL=0 python -m pytest tests/test_gpu_optimizer.py -m gpu -q 2>&1 | tail -1; done
python -m pytest tests/test_gpu_optimizer.py -m gpu -q -x 2>&1 | grep -E "^E " | head -20
