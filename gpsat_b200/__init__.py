"""gpsat_b200 - B200-native local-expert GP engine behind GPSat's BaseGPRModel plugin surface."""
__version__ = "0.1.0"
