"""Host-side mirror of the reference's ``Model.radiation`` package for the real-gas path."""
