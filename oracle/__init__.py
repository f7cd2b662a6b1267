"""TEST INFRASTRUCTURE: CPU oracle for the detectFaces hot path (see haar_oracle.c)."""
