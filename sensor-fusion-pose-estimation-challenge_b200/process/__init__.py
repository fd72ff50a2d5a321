"""Drop-in for the reference package path `process` (modules/lidar/process/)."""
