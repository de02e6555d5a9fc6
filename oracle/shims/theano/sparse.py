"""`theano.sparse` stand-in -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.  SNN_RBM.py:10 imports the module and never uses it."""
