"""gp_dla_detection_b200 -- B200-native drop-in for the DLA model-evaluation hot path of
sbird/gp_dla_detection (process_qsos.m sample loop + voigt.c + log_mvnpdf_low_rank.m).

The compute lives in libdla_b200.so (hand-written sm_100a CUDA behind the C ABI of
include/dla_b200.h); this package is the Python host side that mirrors the reference's
interface.  There is no CPU fallback.
"""
from .set_parameters import DEFAULT as default_parameters, Parameters  # noqa: F401
from .process_qsos import (DLASamples, DlaContext, LearnedModel, PackedBatch, PreparedQuasar,  # noqa: F401
                           dla_existence_prior, finish_outputs, prepare_quasar, process_qsos)

__all__ = ["Parameters", "default_parameters", "DLASamples", "DlaContext", "LearnedModel", "PackedBatch",
           "PreparedQuasar", "dla_existence_prior", "finish_outputs", "prepare_quasar", "process_qsos"]
