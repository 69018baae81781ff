"""predict_hf.py with the B200 scorer behind the same call (reference: predict_hf.py:5-40).

Enter the clinical parameters of one patient below, exactly as in the reference script; the model is read from the flat
blob (``tools/convert_model.py``) or, if a path to ``hf_predict_model.pkl`` is given, statically decoded from it --
nothing is ever unpickled.

    python tools/predict_hf_b200.py [model.blob | hf_predict_model.pkl]
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import mlr_b200 as P  # noqa: E402

Patient_Params = {
    'Obstructive HCM': 1, 'Gender': 1, 'Syncope': 0, 'Dyspnea': 0, 'Fatigue': 1, 'Presyncope': 0, 'NYHA_Class': 1,
    'Atrial_Fibrillation': 1, 'Hypertension': 0, 'Beta_blocker': 0, 'Ca_Channel_Blockers': 0, 'ACEI_ARB': 0,
    'Coumadin': 0, 'Max_Wall_Thick': 13, 'Septal_Anterior_Motion': 0, 'Mitral_Regurgitation': 0,
    'Ejection_Fraction': 55,
}

path = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "tests", "golden", "published_model.blob")
clf_loaded = (P.HFStackB200.from_published_pickle(path) if path.endswith(".pkl") else P.HFStackB200.from_blob_file(path))
input_vars = [Patient_Params[key] for key in Patient_Params.keys()]
score = clf_loaded.predict_proba(np.reshape(input_vars, (1, -1)))

print('###########################################')
print('Probability of progressive HF is: {0:.2f} %'.format(100 * score[0, 1]))
print('###########################################')
