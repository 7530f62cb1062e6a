"""Generates tests/golden/dbow2_transform.npz by running the REFERENCE's DBoW2 (oracle/_ref/libref_dbow2.so, built from
/root/reference/Thirdparty/DBoW2 by oracle/ref_dbow2/Makefile) on small synthetic text vocabularies: per-feature
(word, weight, node) for several levelsup values and the flattened BowVector / FeatureVector for every scoring / weighting pair
the tests use.  Run here (needs /root/reference):  python tools/gen_dbow2_golden.py"""
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from oracle.ref_dbow2 import ref_dbow2 as R  # noqa: E402
from test_vocabulary import GOLDEN_CASES, features, make_vocabulary  # noqa: E402

out = {}
with tempfile.TemporaryDirectory() as tmp:
    for case, (seed, k, L, scoring, weighting) in enumerate(GOLDEN_CASES):
        voc = make_vocabulary(seed, k, L, p_early=0.12, p_stop=0.1, scoring=scoring, weighting=weighting)
        path = os.path.join(tmp, f"voc{case}.txt")
        voc.to_text(path, trailing_newline=False)   # with it the reference reads an uninitialised parent id (UB, crashes here)
        f = features(seed, voc, 400)
        ref = R.RefVocabulary(path)
        out[f"c{case}_text"] = np.frombuffer(open(path, "rb").read(), np.uint8)
        out[f"c{case}_feat"] = f
        out[f"c{case}_size"] = np.array([ref.size()])
        for lu in range(0, L + 2):
            w, wt, nd = ref.transform_features(f, lu)
            out[f"c{case}_lu{lu}_word"], out[f"c{case}_lu{lu}_weight"], out[f"c{case}_lu{lu}_node"] = w, wt, nd
        bow, fv = ref.transform(f, 2)
        out[f"c{case}_bow_word"] = np.array(list(bow.keys()), np.uint32)
        out[f"c{case}_bow_value"] = np.array(list(bow.values()), np.float64)
        out[f"c{case}_fv_node"] = np.array(list(fv.keys()), np.uint32)
        out[f"c{case}_fv_count"] = np.array([len(v) for v in fv.values()], np.int32)
        out[f"c{case}_fv_index"] = np.array([i for v in fv.values() for i in v], np.uint32)
dst = os.path.join(ROOT, "tests", "golden", "dbow2_transform.npz")
np.savez_compressed(dst, **out)
print(dst, os.path.getsize(dst), "bytes")
