"""Reference-compatible import path: `from models.generator.ngswin import NGswin`."""
