"""textblaster_b200 -- B200-native implementation of TextBlaster's per-document quality-filter hot
path (C4QualityFilter, GopherQualityFilter, GopherRepetitionFilter) behind the reference's
ProcessingStep / pipeline_config.yaml / TextDocument boundary.

The compute lives in hand-written sm_100a CUDA kernels exposed through a C ABI
(include/textblaster_b200.h, libtextblaster_b200.so).  This package is the host-side mirror of the
reference interface for that path; it has no CPU fallback.
"""
from .config import (C4QualityParams, GopherQualityParams, GopherRepetitionParams, PipelineConfig, StepConfig,
                     load_pipeline_config)
from .data_model import ProcessingOutcome, TextDocument
from .error import (ConfigError, ConfigValidationError, DocumentFiltered, PipelineError, StepError, Unexpected)

__all__ = [
    "C4QualityParams", "GopherQualityParams", "GopherRepetitionParams", "PipelineConfig", "StepConfig",
    "load_pipeline_config", "ProcessingOutcome", "TextDocument", "ConfigError", "ConfigValidationError",
    "DocumentFiltered", "PipelineError", "StepError", "Unexpected",
]


def __getattr__(name):  # executor pulls in numpy + the native library lazily
    if name in ("ProcessingStep", "PipelineExecutor", "GpuBatchPipeline", "C4QualityFilter", "GopherQualityFilter",
                "GopherRepetitionFilter", "FineWebQualityFilter", "build_pipeline_from_config", "BatchResult"):
        from . import executor
        return getattr(executor, name)
    raise AttributeError(name)
