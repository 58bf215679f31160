# Source me: environment for building and running (the reference's load-env.sh loads site modules
# and sets CUDAARCHS / MPICH_GPU_SUPPORT_ENABLED; none of that is needed here - no MPI).
export TENZING_ACK_NOTICE=1
export CUDA_HOME=${CUDA_HOME:-/usr/local/cuda}
export PATH=$CUDA_HOME/bin:$PATH
export CUDAARCHS=100a
