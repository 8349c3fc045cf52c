"""B200-native exhaustive descriptor matching (hot path of Tarboucha/3DReconstruction)."""
