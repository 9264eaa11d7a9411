"""bench.py's output contract, checked without a GPU on the reference arm: exactly one line on stdout, valid JSON with
the keys the driver reads -- whatever libraries print while the bench runs goes to stderr (bench.OneLineStdout)."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_prints_exactly_one_json_line():
    cmd = [sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0",
           "--cpu-sample-rows", "20000", "--features", "20", "--depth", "4"]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = out.stdout.splitlines()
    assert len(lines) == 1, lines
    line = json.loads(lines[0])
    for key in ("impl", "metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "config",
                "cpu_baseline", "e2e"):
        assert key in line, key
    assert line["impl"] == "reference" and line["value"] > 0 and line["cpu_baseline"]["kind"] in ("port", "reference")
    assert line["e2e"]["h2d_bytes_per_step"] == 0 and line["e2e"]["d2h_bytes_per_step"] == 0


def test_stdout_guard_sends_stray_prints_to_stderr():
    code = ("import os, sys, json\n"
            "sys.path.insert(0, %r)\n"
            "import bench\n"
            "with bench.OneLineStdout() as out:\n"
            "    print('a library banner')\n"
            "    os.write(1, b'printf from C\\n')\n"
            "    out.emit({'ok': 1})\n"
            "    print('after the line')\n" % ROOT)
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=120, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    assert out.stdout.splitlines() == [json.dumps({"ok": 1})]
    assert "a library banner" in out.stderr and "printf from C" in out.stderr and "after the line" in out.stderr
