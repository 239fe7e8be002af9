"""`python main_gather.py ...` -- same command line as the reference's main_gather.py; the work happens in
distributed_pytorch_b200.cli.run("gather")."""
from distributed_pytorch_b200.cli import main

if __name__ == "__main__":
    main("gather")
