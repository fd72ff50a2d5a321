"""Drop-in for the reference package path `common` (modules/lidar/common/)."""
