// empty stand-in for the CUDA samples helper_functions.h
