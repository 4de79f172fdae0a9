"""robustmvd_b200: B200-native plane-sweep cost-volume path behind the rmvd model API."""
