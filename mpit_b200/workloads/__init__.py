"""Workloads: model + data + training loop (reference layer L5: asyncsgd/goot.lua, BiCNN/bicnn.lua)."""
