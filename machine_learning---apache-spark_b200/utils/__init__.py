"""Utilities: device RNG/step counters, CUDA-event timers, logging, checkpoint helpers."""
