"""Data layer: libsvm reader, synthetic generators with the reference datasets' shapes, the text pipeline
(tokenizer / vocab / transforms) equivalents, samplers."""
