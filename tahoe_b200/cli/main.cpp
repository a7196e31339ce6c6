// Tahoe MODEL DATA -- command-line clone of the reference's main.cu:7-97 over the C ABI.
// Same console strings: performance-model pick, per-strategy
// microseconds per sample, verdicts, speed-up over the baseline strategy.
#include <cstdio>

#include "../../include/BaseTahoeTest.h"

int main(int argc, char** argv) {
  if (argc != 3) {
    std::printf("Usage: %s MODEL_FILE DATA_FILE\n", argv[0]);   // the reference prints and carries on (main.cu:10-13)
    return 2;
  }
  std::printf("Model: %s , Data: %s\n", argv[1], argv[2]);
  BaseTahoeTest* pTest = new BaseTahoeTest(argv[1], argv[2]);
  // The reference evaluates its model BEFORE the forest is loaded (main.cu:24-26 vs
  // BaseTahoeTest.h:82); here the pick is made on the loaded forest and batch.
  float speedup = 0.0f;
  const int best = pTest->SetUp(speedup);
  if (best == 0) {
    std::printf("FAIL: no strategy ran\n");
    delete pTest;
    return 1;
  }
  const int model = pTest->model_choice();
  std::printf("Performance model choose #%d strategy.\n", model);
  std::printf(model == best ? "Performance model predicts correctly\n" : "Performance model predicts incorrectly\n");
  std::printf("Tahoe brings %fx speedup.\n", speedup);
  const bool good = pTest->last_compare_ok;
  pTest->Free();
  delete pTest;
  return good ? 0 : 3;
}
