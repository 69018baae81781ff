"""Convert a fitted stack into the flat model blob (replaces shipping a pickle).

    python tools/convert_model.py published <path/to/hf_predict_model.pkl> <out.blob>   # static decode, never unpickled
    python tools/convert_model.py info <model.blob>

A live scikit-learn object is converted in-process: ``mlr_b200.export.write_blob(mlr_b200.export.arrays_from_stack(clf))``.
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import mlr_b200 as P  # noqa: E402


def main(argv):
    if len(argv) == 4 and argv[1] == "published":
        arrays = P.export.arrays_from_published_pickle(argv[2])
        blob = P.export.write_blob(arrays)
        with open(argv[3], "wb") as fh:
            fh.write(blob)
        print(f"wrote {argv[3]}: {len(blob)} bytes, {len(arrays.members)} members")
    elif len(argv) == 3 and argv[1] == "info":
        arrays = P.export.read_blob(open(argv[2], "rb").read())
        for m in arrays.members:
            extra = (f"{m.sv.shape[0]} support vectors" if m.kind == "svc" else
                     f"{len(m.trees)} trees" if m.kind == "gbt" else "17 coefficients")
            print(f"  {m.kind:6s} {extra}")
        print("  combiner", arrays.final_coef.tolist(), arrays.final_intercept)
    else:
        print(__doc__)
        return 2
    return 0


if __name__ == "__main__":
    sys.exit(main(sys.argv))
