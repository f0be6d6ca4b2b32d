"""artis_b200: B200-native ARTIS packet transport (host-side mirror of the reference interface)."""
