"""brp_b200: B200-native Einstein@Home BRP per-template search loop."""
