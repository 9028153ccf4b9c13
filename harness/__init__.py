"""Test / measurement harness around the reference's unchanged tracker (not product code): shims for absent third-party
packages, the synthetic world, the teacher network, the clip runner."""
