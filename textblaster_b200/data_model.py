"""Data model: ``TextDocument`` and ``ProcessingOutcome`` (src/data_model.rs:5-34)."""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, Optional, Tuple


@dataclass
class TextDocument:
    """src/data_model.rs:5-13.  ``added`` / ``created`` are carried through untouched."""

    id: str = ""
    content: str = ""
    source: str = ""
    added: Optional[object] = None
    created: Optional[Tuple[object, object]] = None
    metadata: Dict[str, str] = field(default_factory=dict)


@dataclass
class ProcessingOutcome:
    """src/data_model.rs:19-34: ``Success(doc)`` | ``Filtered{document, reason}`` | ``Error{...}``."""

    kind: str  # "Success" | "Filtered" | "Error"
    document: TextDocument
    reason: Optional[str] = None
    error_message: Optional[str] = None
    worker_id: Optional[str] = None

    @classmethod
    def success(cls, doc):
        return cls("Success", doc)

    @classmethod
    def filtered(cls, doc, reason):
        return cls("Filtered", doc, reason=reason)
