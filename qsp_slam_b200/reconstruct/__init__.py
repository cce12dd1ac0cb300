"""Mirror of the reference's ``reconstruct`` package for the object-reconstruction hot path only."""
