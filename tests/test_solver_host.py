"""Host logic of the scheduler mirror, CPU only: conf parsing the way arg_parser.h + protobuf text
format read it, and the progress-table rows against rows printed by the reference's own Progress structs
(tests/golden/golden_progress.json, made by tests/golden/make_progress_golden.py through oracle/_ref)."""
import io
import json
import os

from wormhole_b200.solver import Config, DifactoProgress, DifactoStop, Progress, match_files

G = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "golden_progress.json")))


def test_linear_progress_rows_equal_the_reference():
    assert Progress.HeadStr() == G["linear"]["head"]
    p = Progress()
    for v, want in zip(G["linear"]["in"], G["linear"]["rows"]):
        p.data = list(v)
        assert p.PrintStr() == want


def test_difacto_progress_rows_equal_the_reference():
    assert DifactoProgress.HeadStr() == G["difacto"]["head"]
    p = DifactoProgress()
    for v, want in zip(G["difacto"]["in"], G["difacto"]["rows"]):
        p.data = list(v)
        assert p.PrintStr() == want


def test_demo_conf_and_argument_overrides():
    demo = '# demo conf\n\ntrain_data = "learn/data/agaricus.txt.train"\nval_data = "learn/data/agaricus.txt.test"\n' \
           'model_out = "./agaricus_model"\nmax_data_pass = 3\n'                 # learn/linear/guide/demo.conf
    c = Config(demo)
    assert c.train_data == "learn/data/agaricus.txt.train" and c.max_data_pass == 3 and c.model_out == "./agaricus_model"
    assert c.algo == "FTRL" and c.minibatch == 1000 and c.rand_shuffle == 10 and c.max_concurrency == 2   # defaults
    assert not c.has("lr_eta") and c.lr_eta == 0.01
    c.merge("lr_eta=.1 algo=ADAGRAD train_data='a=b.txt' minibatch: 500")         # '=' kept inside quotes
    assert c.has("lr_eta") and abs(c.lr_eta - 0.1) < 1e-7 and c.algo == "ADAGRAD" and c.train_data == "a=b.txt"
    assert c.minibatch == 500
    try:
        Config("no_such_key = 1")
        assert False
    except Exception as e:
        assert "unknown field" in str(e)


def test_match_files(tmp_path):
    for n in ("part-0", "part-1", "other"):
        (tmp_path / n).write_text("x")
    assert [os.path.basename(f) for f in match_files(str(tmp_path / "part-.*"))] == ["part-0", "part-1"]
    assert match_files(str(tmp_path / "other")) == [str(tmp_path / "other")]


def test_difacto_stop_rule():
    out = io.StringIO()
    s = DifactoStop(early_stop=True, min_objv_decr=1e-3, max_objv=2.0, out=out)
    assert not s(500.0, 1000, True) and s(2500.0, 1000, True)         # training: objective above max_objv
    assert not s(600.0, 1000, False)                                   # 100 -> 0.6
    assert not s(590.0, 1000, False)                                   # improved by 0.01
    assert s(589.9, 1000, False) and "minimal requirement" in out.getvalue()


def test_cpp_progress_rows_equal_the_reference(tmp_path):
    """the C++ facade's Progress structs (wormhole_b200/host/{linear,difacto}/progress.h) against the same golden rows"""
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = str(tmp_path / "progress_test")
    r = subprocess.run(["/usr/bin/g++", "-std=c++14", "-I" + os.path.join(root, "wormhole_b200", "host"),
                        os.path.join(root, "tests", "cpp", "progress_test.cc"), "-o", exe], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    inp = "".join("L " + " ".join(repr(float(x)) for x in v) + "\n" for v in G["linear"]["in"])
    inp += "".join("D " + " ".join(repr(float(x)) for x in v) + "\n" for v in G["difacto"]["in"])
    out = subprocess.run([exe], input=inp, capture_output=True, text=True).stdout.split("\n")
    assert out[0] == G["linear"]["head"] and out[1] == G["difacto"]["head"]
    n = len(G["linear"]["rows"])
    assert out[2:2 + n] == G["linear"]["rows"]
    assert out[2 + n:2 + n + len(G["difacto"]["rows"])] == G["difacto"]["rows"]


def test_cpp_text_split_tiles_the_file_at_line_starts(tmp_path):
    """TextSplit of the C++ reader (host/base/minibatch_iter.h): the chunks of all parts, in order, are the file, and
    every chunk starts at a line start and ends with a line break -- the same tiling as reader.part_range."""
    import subprocess
    import numpy as np
    from tests.util import criteo_text
    from wormhole_b200.reader import part_range
    from wormhole_b200 import build
    build.build()
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = str(tmp_path / "textsplit_test")
    r = subprocess.run(["/usr/bin/g++", "-std=c++14", "-I" + os.path.join(root, "include"),
                        "-I" + os.path.join(root, "wormhole_b200", "host"), os.path.join(root, "tests", "cpp", "textsplit_test.cc"),
                        "-o", exe, "-L" + os.path.join(root, "wormhole_b200"), "-lwormhole_b200",
                        "-Wl,-rpath," + os.path.join(root, "wormhole_b200"), "-L/usr/local/cuda/lib64", "-lcudart"],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    text = criteo_text(np.random.default_rng(12), 400, crlf=True)
    f = tmp_path / "c.txt"
    f.write_bytes(text)
    for nparts in (1, 4):
        out = subprocess.run([exe, str(f), str(nparts), "5000"], capture_output=True, text=True).stdout.splitlines()
        pos = 0
        for line in out:
            t = line.split()
            if t[1] == "read":
                b, e = part_range(str(f), int(t[0]), nparts)
                assert int(t[2]) == e - b and pos == e
                continue
            size, first, last = int(t[1]), int(t[2]), int(t[3])
            assert text[pos] == first and text[pos + size - 1] == last and last in (10, 13)
            assert pos == 0 or text[pos - 1] in (10, 13)                 # a chunk starts at a line start
            pos += size
        assert pos == len(text)


def test_cpp_config_parses_demo_conf_and_overrides(tmp_path):
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = str(tmp_path / "config_test")
    r = subprocess.run(["/usr/bin/g++", "-std=c++14", "-I" + os.path.join(root, "wormhole_b200", "host"),
                        os.path.join(root, "tests", "cpp", "config_test.cc"), "-o", exe], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    out = subprocess.run([exe], capture_output=True, text=True).stdout.splitlines()
    assert out == ["parse 1", "train learn/data/agaricus.txt.train",
                   "pass 3 minibatch 1000 shuffle 10 conc 2 algo 3 loss 2",      # defaults of config.proto
                   "has_eta 0 eta 0.01", "override 1", "has_eta 1 eta 0.1 algo 2 minibatch 500",
                   "unknown 0", "badenum 0"]
