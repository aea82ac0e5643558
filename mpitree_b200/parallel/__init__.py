"""Multi-rank runtime: communicators, symmetric peer memory, level scheduler, hand-off."""
