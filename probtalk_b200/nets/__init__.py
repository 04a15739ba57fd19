"""Mirror of the reference's ``nets`` package for the hot path."""
