// empty stand-in: nothing from <TF1.h> is used on the hot path
