"""B200-native rule-generation hot path (package init filled in below)."""
