"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference path (see iu_oracle.cpp)."""
