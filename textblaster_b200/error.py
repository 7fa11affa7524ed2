"""Error types: the variants of the reference's ``PipelineError`` (src/error.rs:9-58) that the
filter path can produce."""
from __future__ import annotations


class PipelineError(Exception):
    """Base of all errors (``enum PipelineError``, src/error.rs:9-10)."""

    def __init__(self, message: str = ""):
        super().__init__(message)
        self.message = message


class ConfigError(PipelineError):
    """``PipelineError::ConfigError`` -- "Configuration error: {0}" (src/error.rs:11-12)."""

    def __str__(self):
        return f"Configuration error: {self.message}"


class ConfigValidationError(PipelineError):
    """``PipelineError::ConfigValidationError`` -- "Configuration validation error: {0}" (src/error.rs:62-63)."""

    def __str__(self):
        return f"Configuration validation error: {self.message}"


class DocumentFiltered(PipelineError):
    """``PipelineError::DocumentFiltered { document, reason }`` (src/error.rs:33-37): the filtered
    document travels inside the error, with the metadata the step wrote."""

    def __init__(self, document, reason: str):
        super().__init__(reason)
        self.document = document
        self.reason = reason

    def __str__(self):
        return f"Document '{self.document.id}' filtered out: {self.reason}"


class StepError(PipelineError):
    """``PipelineError::StepError { step_name, source }`` (src/error.rs:39-43): how the executor
    wraps any error a step returns (src/executor.rs:37-46)."""

    def __init__(self, step_name: str, source: PipelineError):
        super().__init__(str(source))
        self.step_name = step_name
        self.source = source

    def __str__(self):
        return f"Error in processing step '{self.step_name}': {self.source}"


class Unexpected(PipelineError):
    """``PipelineError::Unexpected`` (src/error.rs:65-66); also raised for CUDA / library failures."""

    def __str__(self):
        return f"Unexpected error: {self.message}"
