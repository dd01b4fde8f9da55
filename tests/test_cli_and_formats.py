"""The cudapoa command line tool and the window file formats, mirroring the behaviour of
the reference's tool (cudapoa/src/main.cpp:129-293, application_parameters.cpp:29-150, utils.hpp:60-157).
CPU: option validation and the parsers (compiled into a small host program).  GPU: the tool's stdout against the
library driven through the C ABI on the same windows."""
import os
import subprocess
import sys
import textwrap

import pytest

import parity
from claragenomicsanalysis_b200 import build
from claragenomicsanalysis_b200.windows import generate_windows

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def cli():
    build.build()
    assert os.path.exists(build.CLI)
    return build.CLI


def write_cudapoa_file(path, ws):
    with open(path, "w") as f:
        for w in range(len(ws)):
            reads = ws.window(w)
            f.write("%d\n" % len(reads))
            for r in reads:
                f.write(r + "\n")


def write_fasta(path, reads, width=60):
    with open(path, "w") as f:
        for i, r in enumerate(reads):
            f.write(">read_%d some description\n" % i)
            for a in range(0, len(r), width):
                f.write(r[a:a + width] + "\n")


def run_cli(cli, *args):
    return subprocess.run([cli] + list(args), stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=600)


# ---- CPU ----

def test_cli_rejects_bad_options(cli, tmp_path):
    good = tmp_path / "w.txt"
    good.write_text("2\nACGT\nACGT\n")
    assert run_cli(cli).returncode == 1  # no input: usage
    r = run_cli(cli, "-i", str(tmp_path / "missing.txt"))
    assert r.returncode == 1 and "Invalid input file" in r.stderr
    for bad in (["-R", "1.5"], ["-R", "0"], ["-m", "-1"], ["-n", "3"], ["-g", "2"], ["-M", "0"], ["-b", "0"],
                ["-A", "-f"]):
        r = run_cli(cli, "-i", str(good), *bad)
        assert r.returncode == 1, bad
    other = tmp_path / "w2.txt"
    other.write_text("1\nACGT\n")
    r = run_cli(cli, "-i", str(good), "-i", str(other))  # two cudapoa-format files
    assert r.returncode == 1 and "Invalid input" in r.stderr
    assert run_cli(cli, "-h").returncode == 0


def test_window_file_parsers(tmp_path):
    """parse_cudapoa_file / parse_fasta_files / resize_windows of include/.../cudapoa/utils.hpp."""
    src = tmp_path / "p.cpp"
    src.write_text(textwrap.dedent(r'''
        #include <claraparabricks/genomeworks/cudapoa/utils.hpp>
        #include <iostream>
        using namespace claraparabricks::genomeworks::cudapoa;
        static void dump(const std::vector<std::vector<std::string>>& w) {
            for (auto& g : w) { std::cout << g.size(); for (auto& s : g) std::cout << ' ' << s; std::cout << '\n'; }
        }
        int main(int argc, char** argv) {
            std::vector<std::vector<std::string>> a, b, c;
            parse_cudapoa_file(a, argv[1], -1); dump(a); std::cout << "--\n";
            parse_cudapoa_file(b, argv[1], 5); dump(b); std::cout << "--\n";
            parse_fasta_files(c, {argv[2], argv[3]}, 1); dump(c);
            return 0;
        }'''))
    (tmp_path / "w.txt").write_text("2\nACGT\nACGA\n3\nTT\nTTA\nTA\n")
    (tmp_path / "a.fa").write_text(">r0 x\nACG\nT\n>r1\nGG\n")
    (tmp_path / "b.fa").write_text(">q\nAAAA\n")
    exe = tmp_path / "p"
    subprocess.check_call(["g++", "-std=c++17", "-I" + os.path.join(ROOT, "include"), "-I/usr/local/cuda/include",
                           str(src), "-o", str(exe)])
    out = subprocess.check_output([str(exe), str(tmp_path / "w.txt"), str(tmp_path / "a.fa"), str(tmp_path / "b.fa")],
                                  text=True)
    all_w, five, fasta = out.split("--\n")
    assert all_w == "2 ACGT ACGA\n3 TT TTA TA\n"
    assert five == "2 ACGT ACGA\n3 TT TTA TA\n2 ACGT ACGA\n3 TT TTA TA\n2 ACGT ACGA\n"  # cyclic repetition
    assert fasta == "2 ACGT GG\n"  # one window per FASTA file, truncated to one group


# ---- GPU ----

@pytest.mark.gpu
def test_cli_consensus_msa_dot_match_library(cli, tmp_path):
    from claragenomicsanalysis_b200 import _cabi

    lib = _cabi.native()
    ws = generate_windows(10, 31000, template_length=300, n_reads=8)
    path = tmp_path / "windows.txt"
    write_cudapoa_file(path, ws)
    want = parity.run_abi(lib, ws, 512, 8, band=256, banded=True, output="consensus")
    dot = tmp_path / "g.dot"
    r = run_cli(cli, "-i", str(path), "-d", str(dot))
    assert r.returncode == 0, r.stderr
    assert r.stdout.split("\n")[:-1] == want["consensus"]
    assert "Processed groups 0 - 9 (batch 0)" in r.stderr
    assert dot.read_text().count("digraph") == len(ws)

    want_msa = parity.run_abi(lib, ws, 512, 8, band=256, banded=True, output="msa")
    r = run_cli(cli, "-i", str(path), "-a")
    assert r.returncode == 0, r.stderr
    assert r.stdout.split("\n")[:-1] == [row for g in want_msa["msa"] for row in g]

    want_full = parity.run_abi(lib, ws, 512, 8, banded=False, output="consensus")
    r = run_cli(cli, "-i", str(path), "-f", "-M", "14")  # full alignment, groups repeated up to 14
    assert r.returncode == 0, r.stderr
    assert r.stdout.split("\n")[:-1] == (want_full["consensus"] * 2)[:14]

    r = run_cli(cli, "-i", str(path), "-A", "-b", "128")
    want_ad = parity.run_abi(lib, ws, 512, 8, band=128, banded="adaptive", output="consensus")
    assert r.returncode == 0 and r.stdout.split("\n")[:-1] == want_ad["consensus"]


@pytest.mark.gpu
def test_cli_fasta_inputs(cli, tmp_path):
    from claragenomicsanalysis_b200 import _cabi

    lib = _cabi.native()
    ws = generate_windows(3, 32000, template_length=200, n_reads=6)
    paths = []
    for w in range(len(ws)):
        p = tmp_path / ("w%d.fasta" % w)
        write_fasta(p, ws.window(w))
        paths.append(str(p))
    args = []
    for p in paths:
        args += ["-i", p]
    r = run_cli(cli, *args, "-m", "5", "-n", "-4", "-g", "-8")
    assert r.returncode == 0, r.stderr
    want = parity.run_abi(lib, ws, 256, 6, band=256, banded=True, output="consensus", gap=-8, mismatch=-4, match=5)
    assert r.stdout.split("\n")[:-1] == want["consensus"]
