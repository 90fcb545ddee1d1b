#!/usr/bin/env python3
"""Copies the reference's shipped binary fixtures for the find_mems path into tests/golden/ and records
their SHA-256 (run in the build container, where /root/reference exists; the GPU box only sees the copies).

These are DATA files the reference ships under test_data/ (texts, grlBWT .rl_bwt, FastLocate .ri, tags);
no reference source code is copied. SURVEY.md App. D lists what each one is.
Usage: python tests/golden/make_golden.py
"""
import hashlib, json, os, shutil, sys
REF = "/root/reference/test_data"
HERE = os.path.dirname(os.path.abspath(__file__))
FILES = {
    "xy.ri": "bidirectional_test/xy.ri",
    "xy_bidirectional_compressed.tags": "bidirectional_test/xy_bidirectional_compressed.tags",
    "xy_bidirectional.tags": "bidirectional_test/xy_bidirectional.tags",
    "contigs_xy": "bidirectional_test/contigs_xy",
    "contigs_xy.rl_bwt": "bidirectional_test/contigs_xy.rl_bwt",
    "contigs_x": "bidirectional_test/contigs_x",
    "contigs_x.rl_bwt": "bidirectional_test/contigs_x.rl_bwt",
    "reads.txt": "bidirectional_test/reads.txt",
    "small_test.txt": "bidirectional_test/small_test/test.txt",
    "small_test.rl_bwt": "bidirectional_test/small_test/test.rl_bwt",
    "two_contig_xy.ri": "two_contig_graph/r-index/xy.ri",
    "contigs_XY.txt": "two_contig_graph/contigs_XY.txt",
    "contigs_XY.rl_bwt": "two_contig_graph/contigs_XY.rl_bwt",
    "x.newline_separated": "x.newline_separated",
    "x.rl_bwt": "x.rl_bwt",
    "med_test.txt": "med_test.txt",
    "med_test.rl_bwt": "med_test.rl_bwt",
}
def main():
    if not os.path.isdir(REF):
        sys.exit("reference test_data not found (this script only runs in the build container)")
    manifest = {}
    for dst, src in sorted(FILES.items()):
        p = os.path.join(REF, src)
        shutil.copyfile(p, os.path.join(HERE, dst))
        os.chmod(os.path.join(HERE, dst), 0o644)
        manifest[dst] = {"source": "test_data/" + src, "bytes": os.path.getsize(p),
                         "sha256": hashlib.sha256(open(p, "rb").read()).hexdigest()}
    json.dump(manifest, open(os.path.join(HERE, "MANIFEST.json"), "w"), indent=1, sort_keys=True)
    print("copied", len(manifest), "fixtures")
if __name__ == "__main__":
    main()
