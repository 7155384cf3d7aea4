"""Small cross-cutting helpers (phase tracing / profiling hooks)."""
from .phases import PhaseTrace  # noqa: F401
