"""TEST INFRASTRUCTURE ONLY — CPU restatements of the reference's algorithms for the hot path.
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import anything from here; the product (soar_b200/, diff_gaussian_rasterization/, simple_knn/)
never does and fails loudly when its CUDA library is missing."""
