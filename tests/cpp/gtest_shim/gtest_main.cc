// gtest_main stand-in (see gtest/gtest.h beside this file).
#include "gtest/gtest.h"
int main(int argc, char** argv) {
  ::testing::InitGoogleTest(&argc, argv);
  return RUN_ALL_TESTS();
}
