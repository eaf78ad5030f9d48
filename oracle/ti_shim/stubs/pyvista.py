"""Empty stand-in so that the reference object.py:8-10 imports (test infrastructure; mesh ingestion is injected)."""
