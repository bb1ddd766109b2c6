"""B200-native gain-map hot path (generate / apply / toneMap) behind the libultrahdr API."""
