from .gps import GPSConv
