timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "bvh_build or gpu_semantics" 2>&1 | tail -2
bash tools/gpu_cmd2.sh | grep -E "collapse|refit"
python tools/profile_frame.py world16 3840 2160 9 4 | tail -1 | cut -c1-400
