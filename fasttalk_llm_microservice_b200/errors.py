"""The error convention of the handler boundary.

Inside the reference process the handler raises the reference's own
`LLMServiceError` (`app/utils/error_handler.py:50-76`) so `WebSocketLLMServer` special-cases it
(`websocket_server_vllm.py:438-444`).  Outside it (tests, bench, the GPU box) an identically
shaped stand-in is used.
"""
from __future__ import annotations

from enum import Enum
from typing import Any, Dict, Optional

try:  # running inside the reference service: use its types
    from app.utils.error_handler import ErrorCategory, ErrorSeverity, LLMServiceError  # type: ignore
except Exception:  # noqa: BLE001 - reference not importable here

    class ErrorCategory(Enum):
        CONNECTION = "connection"
        PROCESSING = "processing"
        RESOURCE = "resource"
        CONFIGURATION = "configuration"
        SYSTEM = "system"
        GPU = "gpu"
        TIMEOUT = "timeout"
        VALIDATION = "validation"

    class ErrorSeverity(Enum):
        LOW = "low"
        MEDIUM = "medium"
        HIGH = "high"
        CRITICAL = "critical"

    class LLMServiceError(Exception):
        def __init__(self, message: str, category: ErrorCategory = ErrorCategory.SYSTEM,
                     severity: ErrorSeverity = ErrorSeverity.MEDIUM, recoverable: bool = True,
                     retry_after: Optional[float] = None):
            super().__init__(message)
            self.message = message
            self.category = category
            self.severity = severity
            self.recoverable = recoverable
            self.retry_after = retry_after

        def to_dict(self) -> Dict[str, Any]:
            return {"message": self.message, "category": self.category.value,
                    "severity": self.severity.value, "recoverable": self.recoverable,
                    "retry_after": self.retry_after}
