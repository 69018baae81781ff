"""B200-native scorer for the heart-failure-progression stack (one hot path: batched predict_proba / predict).

The directory name carries a hyphen, so import it with ``importlib.import_module("machine-learning-replications_b200")``
or through the ``mlr_b200`` alias module at the repository root.
"""
from .schema import N_FEATURES, load_cohort_mat, patient_vector  # noqa: F401
from . import export, shard, synth, synth_models  # noqa: F401
from . import capi  # noqa: F401,E402
from .facade import HFStackB200, synth_rows_device  # noqa: F401,E402
from . import facade, impute, inspection, metrics  # noqa: F401,E402
from .impute import KNNImputerB200  # noqa: F401,E402
