// TEST INFRASTRUCTURE ONLY -- empty stand-in for ros/package.h
