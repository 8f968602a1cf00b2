"""B200-native offer matching for dstack: jobs x offers predicate + price ranking on sm_100a."""

__version__ = "0.1.0"
