"""Host-side mirror of the reference operator interface (stylegan2_ada/torch_utils/ops/*, training/networks.py)."""
