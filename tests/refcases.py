"""The cases on which oracle/ (the restatement), the CUDA path and the REFERENCE'S OWN CODE (oracle/_ref/libpgiref.so)
are compared: indexes, reads and parameters, all reproducible from seeds and the shipped fixtures in tests/golden/.

Used by tests/golden/make_ref_golden.py (runs the reference's code where /root/reference is mounted and commits what it
printed as tests/golden/ref_find_mems.json), tests/test_ref_pin.py (CPU) and tests/test_gpu_parity.py (GPU)."""
import functools
import numpy as np
import datasets, orc
from conftest import load_package

@functools.lru_cache(maxsize=None)
def x_index():
    """BASELINE config 1: encoded sigma = 5 index rebuilt from the shipped x.rl_bwt (forward-only text, 3 sequences) +
    synthetic aligned sdsl-compact tags: one node per 32 text bases of every sequence, endmarker rows carry 0
    (no tags for x ship with the reference)."""
    prep = load_package().prep
    rl = datasets.gold("x.rl_bwt")
    ri, info, rows = prep.rindex_from_rlbwt(rl, encoded=True, with_samples=True, want_rows=True)
    ml = info["max_length"]; text = datasets.gold("x.newline_separated")
    lens = np.array([len(l) for l in text.split(b"\n") if l])
    seq = (rows // np.uint64(ml)).astype(np.int64); off = (rows % np.uint64(ml)).astype(np.int64)
    val = (((seq * 64 + off // 32 + 1).astype(np.uint64)) << np.uint64(11)) | (off % 32).astype(np.uint64)
    val[off >= lens[seq]] = 0
    tags0, _ = prep.tags_from_rows(val, 0)
    return dict(ri=ri, tags=tags0, layout=0, text=text, info=info)

@functools.lru_cache(maxsize=None)
def xy_compact_tags():
    """Our aligned sdsl-compact tags for the shipped xy.ri (the layout src/find_mems.cpp itself loads)."""
    prep = load_package().prep
    ri, info, rows = prep.rindex_from_rlbwt(datasets.gold("contigs_xy.rl_bwt"), encoded=False, with_samples=True, want_rows=True)
    assert ri == datasets.gold("xy.ri")
    ml = info["max_length"]; text = datasets.gold("contigs_xy")
    lens = np.array([len(l) for l in text.split(b"\n") if l])
    seq = (rows // np.uint64(ml)).astype(np.int64); off = (rows % np.uint64(ml)).astype(np.int64)
    val = (((seq // 2 * 64 + off // 32 + 1).astype(np.uint64)) << np.uint64(11)) | ((seq % 2).astype(np.uint64) << np.uint64(10)) | (off % 32).astype(np.uint64)
    val[off >= lens[seq]] = 0
    return prep.tags_from_rows(val, 0)[0]

def indexes():
    xy_text = datasets.gold("contigs_xy")
    out = {
        "xy_legacy_bytecode": dict(ri=datasets.gold("xy.ri"), tags=datasets.gold("xy_bidirectional_compressed.tags"), layout=1, text=xy_text),
        "xy_legacy_compact": dict(ri=datasets.gold("xy.ri"), tags=xy_compact_tags(), layout=0, text=xy_text),
        "two_contig_legacy": dict(ri=datasets.gold("two_contig_xy.ri"), tags=None, layout=0, text=datasets.gold("contigs_XY.txt")),
        "x_encoded_sigma5": x_index(),
    }
    for enc in (True, False):
        s = datasets.synth_index(encoded=enc)
        out["synth_%s_compact" % ("enc" if enc else "leg")] = dict(ri=s["ri"], tags=s["tags0"], layout=0, text=s["text"])
    s = datasets.synth_index(encoded=True)
    out["synth_enc_bytecode"] = dict(ri=s["ri"], tags=s["tags1"], layout=1, text=s["text"])
    return out

# (index, reads recipe, min_len, min_occ). Recipes: ("shipped",) = tests/golden/reads.txt; ("sample", n, len, sub, seed);
# ("ragged", seed) = empty / short / N / lower-case / NUL / high-byte reads of ragged length (no newline inside a read).
CASES = [
    ("xy_legacy_bytecode", ("shipped",), 5, 1),
    ("xy_legacy_bytecode", ("shipped",), 10, 1),
    ("xy_legacy_bytecode", ("sample", 2500, 150, 0.01, 2), 20, 1),
    ("xy_legacy_bytecode", ("sample", 600, 150, 0.01, 3), 5, 1),
    ("xy_legacy_bytecode", ("ragged", 5), 3, 1),
    ("xy_legacy_compact", ("sample", 2000, 150, 0.01, 4), 20, 1),
    ("xy_legacy_compact", ("sample", 500, 100, 0.02, 5), 10, 2),
    ("two_contig_legacy", ("sample", 1500, 150, 0.01, 6), 20, 1),
    ("x_encoded_sigma5", ("sample", 2500, 150, 0.0, 1), 20, 1),
    ("x_encoded_sigma5", ("sample", 500, 150, 0.0, 1), 5, 1),
    ("synth_enc_compact", ("sample", 2500, 150, 0.01, 7), 20, 1),
    ("synth_enc_compact", ("sample", 500, 150, 0.03, 8), 1, 1),
    ("synth_enc_compact", ("ragged", 9), 20, 1),
    ("synth_enc_compact", ("ragged", 9), 1, 2),
    ("synth_leg_compact", ("sample", 1500, 150, 0.01, 10), 31, 2),
    ("synth_enc_bytecode", ("sample", 1500, 150, 0.01, 11), 10, 3),
]

def reads_of(ix, recipe):
    """-> list of bytes (one per read; the reference's loop skips empty lines, so empty reads are dropped here)."""
    if recipe[0] == "shipped":
        rs = datasets.gold("reads.txt").split(b"\n")
    elif recipe[0] == "sample":
        _, n, L, sub, seed = recipe
        buf, off = datasets.sample_reads(ix["text"], n, L, sub, seed=seed)
        rs = [buf[int(off[i]):int(off[i + 1])].tobytes() for i in range(n)]
    else:
        rs = [r for r in datasets.ragged_reads(ix["text"], seed=recipe[1]) if b"\n" not in r]
    # a byte >= 0x80 is a negative `char` that the reference converts to a huge size_t and indexes complement_table with
    # (algorithm.hpp:667 -> r-index.cpp:726): out of bounds there, so such reads are not part of the pin
    return [r for r in rs if r and max(r) < 0x80]

def case_key(c):
    return "%s|%s|%d|%d" % (c[0], ",".join(str(x) for x in c[1]), c[2], c[3])
