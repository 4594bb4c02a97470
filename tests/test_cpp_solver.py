"""tools/linear_b200.cc -- the reference's `linear.dmlc <conf> key=value ...` entry on top of the C++ facade
(host/solver/minibatch_solver.h: MinibatchScheduler::Run, host/base/minibatch_iter.h, AsgdServer / AsgdWorker): compiled
with g++ against the C-ABI library. CPU: it compiles and links. GPU: the demo.conf job on agaricus text prints the
same console output, progress rows included, as the Python mirror (wormhole_b200/solver.py, itself checked against the
CPU checker in tests/test_gpu_solver.py) run with the same rand() seed, and writes the same model file."""
import ctypes
import io
import os
import re
import subprocess

import pytest

from tests.util import agaricus_text

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def build_binary(tmp):
    from wormhole_b200 import build
    build.build()
    exe = os.path.join(str(tmp), "linear_b200")
    cmd = ["/usr/bin/g++", "-std=c++14", "-O1", "-Wall", "-I" + os.path.join(ROOT, "include"),
           "-I" + os.path.join(ROOT, "wormhole_b200", "host"), os.path.join(ROOT, "tools", "linear_b200.cc"),
           "-o", exe, "-L" + os.path.join(ROOT, "wormhole_b200"), "-lwormhole_b200",
           "-Wl,-rpath," + os.path.join(ROOT, "wormhole_b200"), "-L/usr/local/cuda/lib64", "-lcudart"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return exe


def test_linear_b200_compiles_and_links(tmp_path):
    assert os.path.exists(build_binary(tmp_path))


def strip_seconds(log):
    """progress rows start with the elapsed seconds: drop that column"""
    return [re.sub(r"^\s*\d+  ", "", l) if re.match(r"^\s*\d+  \s*[\d.e+]+", l) else l for l in log.splitlines()]


@pytest.mark.gpu
def test_linear_b200_prints_what_the_python_mirror_prints(tmp_path):
    from wormhole_b200.solver import Config, MinibatchScheduler
    exe = build_binary(tmp_path)
    text = agaricus_text()
    lines = text.split(b"\n")
    train, val = tmp_path / "agaricus.txt.train", tmp_path / "agaricus.txt.test"
    train.write_bytes(text)
    val.write_bytes(b"\n".join(lines[:1500]) + b"\n")
    conf = tmp_path / "demo.conf"
    conf.write_text('# demo conf\n\ntrain_data = "%s"\nval_data = "%s"\nmodel_out = "%s"\nmax_data_pass = 2\n'
                    % (train, val, tmp_path / "cpp_model"))
    args = ["print_sec=1000", "num_parts_per_file=3", "lr_eta=.1"]
    r = subprocess.run([exe, str(conf)] + args + ["-seed", "17"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    # the Python mirror, same conf and overrides, same rand() seed
    pconf = Config.from_file(str(conf), args)
    pconf.model_out = str(tmp_path / "py_model")
    out = io.StringIO()
    ctypes.CDLL("libc.so.6").srand(17)
    assert MinibatchScheduler(pconf, out=out).Run()
    assert strip_seconds(r.stdout) == strip_seconds(out.getvalue())
    a = open(str(tmp_path / "cpp_model_part-0"), "rb").read()
    b = open(str(tmp_path / "py_model_part-0"), "rb").read()
    assert len(a) > 0 and a == b
