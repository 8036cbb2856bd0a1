"""Stand-in for geopy 2.3.0: only ``geopy.distance.distance`` (reference coordsim/reader/reader.py:11,223)."""
