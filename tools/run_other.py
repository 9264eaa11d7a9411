"""Developer tool: one of bench.py's other_configs on its own (single GPU): python tools/run_other.py C1"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import bench  # noqa: E402

torch.cuda.set_device(0)
print(json.dumps(bench.measure_config(sys.argv[1], 1, 0, 0, None, torch.cuda.synchronize)))
