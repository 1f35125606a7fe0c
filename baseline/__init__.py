"""Reference baseline (the UNMODIFIED reference shipped to the GPU box): tests / tools / bench secondary only."""
