"""Empty `matplotlib` stand-in -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.  data_fm.py:1 imports matplotlib.pyplot at module level and
never uses it on the path run here (SNN_DAE.py imports data_fm); matplotlib is not installed in this image, so the import is
satisfied by this empty package.  Nothing is plotted anywhere."""
