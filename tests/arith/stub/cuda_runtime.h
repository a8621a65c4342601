// empty stand-in: the arithmetic check compiles the device headers as host code
