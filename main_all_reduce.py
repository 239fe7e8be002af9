"""`python main_all_reduce.py ...` -- same command line as the reference's main_all_reduce.py; the work happens in
distributed_pytorch_b200.cli.run("all_reduce")."""
from distributed_pytorch_b200.cli import main

if __name__ == "__main__":
    main("all_reduce")
