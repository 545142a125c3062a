"""B200-native (sm_100a) embedding-extraction and trial-scoring path, drop-in for the
inference hot path of hiimmuc/SpeakerVerification (see DESIGN.md / INTEGRATION.md)."""
__version__ = "0.1.0"
