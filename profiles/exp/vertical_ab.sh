python profiles/exp/vertical_one.py 1 4 2>&1 | tail -3
python profiles/exp/vertical_one.py 0 4 2>&1 | tail -3
python -m pytest tests/test_gpu_vertical_forward.py -q -m gpu -x 2>&1 | tail -3
