#!/usr/bin/env python3
"""Build the bench index, dump its .ri and run tools/guess_check.cpp on it (CPU check of the device directory)."""
import os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench
args = bench.parse_args()
w = bench.make_workload(args, 0, 0)
open("/tmp/bench.ri", "wb").write(w["ri"])
subprocess.check_call(["g++", "-O2", "-std=c++17", "-o", "/tmp/guess_check", os.path.join(ROOT, "tools", "guess_check.cpp")], cwd=os.path.join(ROOT, "tools"))
subprocess.check_call(["/tmp/guess_check", "/tmp/bench.ri", "5"])
