"""cubesphere -- host side of the B200-native cubed-sphere shallow-water step."""
