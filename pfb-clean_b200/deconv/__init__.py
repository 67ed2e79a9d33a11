"""Drop-in for pfb/deconv: the CLEAN minor loops (clark, hogbom) on the device."""
