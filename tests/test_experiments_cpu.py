"""Experiment harness, the parts that need no GPU: the figure writer and the sweep scripts."""
import json
import os
import subprocess
import sys
import xml.etree.ElementTree as ET

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXP = os.path.join(ROOT, "experiments")


def test_plot_perf_writes_wellformed_svg(tmp_path):
    rows = []
    for algo, base in (("nocc", 1600.0), ("atomic", 2000.0), ("radix", 2600.0), ("adaptive", 2100.0)):
        for lw in range(0, 28, 3):
            rows.append({"algo": algo, "dataDistr": "local_shuffle", "shuffleRange": 1 << lw,
                         "hashBuildTimeInMicroseconds": base * (1.0 + 0.1 * lw)})
    inp = tmp_path / "w.jsonl"
    inp.write_text("\n".join(json.dumps(r) for r in rows) + "\n")
    out = tmp_path / "perf.svg"
    subprocess.run([sys.executable, os.path.join(EXP, "plot_perf.py"), "--in", str(inp), "--out", str(out)], check=True)
    root = ET.parse(out).getroot()
    ns = "{http://www.w3.org/2000/svg}"
    published = json.load(open(os.path.join(EXP, "published_cpu.json")))["build_probe"]
    n_cpu = sum(1 for k in published if not k.startswith("_") and k != "log2_W")
    assert len(root.findall(ns + "path")) == 4 + n_cpu  # one polyline per curve


def test_published_cpu_numbers_are_the_survey_tables():
    d = json.load(open(os.path.join(EXP, "published_cpu.json")))
    for fam in ("build_probe", "build_only"):
        xs = d[fam]["log2_W"]
        for k, v in d[fam].items():
            if not k.startswith("_") and k != "log2_W":
                assert len(v) == len(xs) and all(x > 0 for x in v)
    # spot values of the reference logs
    assert d["build_only"]["NoCC (CPU)"][0] == 118971 and d["build_probe"]["Atomic (CPU)"][-1] == 2259926


def test_sweep_scripts_parse_and_name_the_reference_sweeps():
    for f in ("sweeps.sh", "runner.sh"):
        subprocess.run(["bash", "-n", os.path.join(EXP, f)], check=True)
    r = subprocess.run(["bash", os.path.join(EXP, "sweeps.sh")], capture_output=True, text=True)
    assert r.returncode == 1
    for name in ("motivation", "SizeToAbortsAndTimeSorted", "SizeToAbortsAndTimeShuffled", "TSizeAndShuffleWindowstoTime",
                 "AtomicsVsHTMVsNoCC", "adaptive", "probe"):
        assert name in r.stdout
