"""Writes tests/golden/kats.json.

Sources of the values (nothing here is computed by the code under test):
  * generators: SURVEY.md 8c "Generator KATs", probed from the reference's
    include/DataGen.hpp compiled with g++ 13 / glibc rand();
  * reference_logs_2^27: deterministic fields of the committed experiment logs,
    experiments/new_backup/AtomicsVsHTMVsNoCC_log1:1-6, probe_log1:1,29,57,
    motivation_log1:8 (PRO Results), npo_probe_log1:8;
  * ref_serial_uniform_2^20: the reference's AtomicHashBuild run serially on the
    "uniform" input (probed).
Run:  python tests/golden/make_kats.py
"""
import json
import os

KATS = {
    "generators": {
        "sorted_2^20": {"dist": "sorted", "n": 1 << 20, "range": 16, "sum": 549756338176,
                        "possum": 384307717958270976},
        "shuffle_2^20": {"dist": "shuffle", "n": 1 << 20, "range": 16, "possum": 288288018717812865,
                         "first8": [488349, 25720, 365654, 670898, 712420, 938430, 1027069, 554635]},
        "local_shuffle_2^20_w16": {"dist": "local_shuffle", "n": 1 << 20, "range": 16,
                                   "possum": 384307717919436259, "first8": [8, 1, 12, 7, 6, 5, 4, 2]},
        "local_shuffle_2^20_w4096": {"dist": "local_shuffle", "n": 1 << 20, "range": 4096,
                                     "possum": 384305006198352694},
        "uniform_2^20": {"dist": "uniform", "n": 1 << 20, "range": 16, "sum": 549507039110,
                         "possum": 384120649505443004, "first8": [10, 3, 4, 2, 14, 3, 3, 3]},
        "shuffle_2^24": {"dist": "shuffle", "n": 1 << 24, "range": 16, "sum": 140737496743936,
                         "possum": 18087968770137856679},
        "local_shuffle_2^24_w65536": {"dist": "local_shuffle", "n": 1 << 24, "range": 65536,
                                      "possum": 6137961876333909458},
        "uniform_2^24": {"dist": "uniform", "n": 1 << 24, "range": 16, "sum": 140719426506879,
                         "possum": 6059732478467025691},
    },
    "reference_logs_2^27": {
        "inputSum": 9007199321849856, "totalMatches": 134217728, "conflicts": 0,
        "nocc_outputSum": 9007199187632128, "PRO_Results_b14": 549688705024, "NPO_Results": 134217728,
    },
    "ref_serial_uniform_2^20": {"conflicts": 176864, "totalMatches": 871712},
    "ref_serial_uniform_2^24": {"conflicts": 2825912},
    "pro_checksums": {"n2^24_b7": 1099503239168, "n2^24_b10": 137430564864, "n2^24_b14": 8581545984,
                      "n2^20_b14": 33030144},
}

if __name__ == "__main__":
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "kats.json")
    with open(path, "w") as f:
        json.dump(KATS, f, indent=1, sort_keys=True)
    print("wrote", path)
