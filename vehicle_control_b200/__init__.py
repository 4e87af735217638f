"""vehicle_control_b200 -- B200-native batched NMPC engine behind the vehicle-control host surface."""
__version__ = "0.1.0"
