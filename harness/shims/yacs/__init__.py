"""Stand-in for the `yacs` package (absent in this image): just enough of CfgNode for dpvo/config.py and the drivers."""
