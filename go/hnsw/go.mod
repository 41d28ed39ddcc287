module github.com/TFMV/hnsw

go 1.24
