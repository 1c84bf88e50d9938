// TEST INFRASTRUCTURE ONLY -- empty stand-in for ros/ros.h
