// ORACLE -- TEST INFRASTRUCTURE ONLY.  (point-cloud preprocessing rows f1/f2; filled in below)
